mkdir -p gpurun_out
python tools/profile_forward.py > gpurun_out/prof_fwd.log 2>&1
python bench.py --backbone f16x3 --steps 5 --warmup 3 --loftr-pairs 0 > gpurun_out/bench_f16x3.json 2> gpurun_out/bench_f16x3.err
python -m pytest tests/test_gpu_conv.py tests/test_gpu_pipeline.py -q 2>&1 | tail -5
tail -70 gpurun_out/prof_fwd.log; tail -c 300 gpurun_out/bench_f16x3.err
