mkdir -p gpurun_out
for s in 1; do echo "lim_scale $s"; OPB200_SCREEN_LIM_SCALE=$s python tools/screen_probe.py 2>&1 | grep '"max_delta": 0.6\|"algo"'; done > gpurun_out/limscale.log 2>&1
python -m pytest tests -m gpu -q 2>&1 | tail -120 > gpurun_out/t2.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench2.json 2> gpurun_out/bench2.err
cat gpurun_out/limscale.log; tail -60 gpurun_out/t2.log; tail -c 300 gpurun_out/bench2.err
