#!/bin/bash
# One GPU pass: parity suite, smoke, bench (both arms), ncu launch list and a full-set capture of the coarse-matching kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r02_gpu.txt
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r02_gputests.log 2>&1
tail -5 gpurun_out/r02_gputests.log
( time timeout 300 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r02_smoke.log 2>&1
tail -8 gpurun_out/r02_smoke.log
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench.json ) 2> gpurun_out/r02_bench.err
tail -c 600 gpurun_out/r02_bench.err; head -c 1500 gpurun_out/r02_bench.json; echo
( time timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_reference.json ) 2> gpurun_out/r02_bench_reference.err
tail -c 300 gpurun_out/r02_bench_reference.err; head -c 600 gpurun_out/r02_bench_reference.json; echo
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_bench_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 > gpurun_out/r02_ncu_launch.log 2>&1
echo launches rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'sim_fast_kernel|merge_fast|pack_f16|cand_' --launch-skip 40 -c 12 -o gpurun_out/r02_coarse_full -f \
  python bench.py --steps 2 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 > gpurun_out/r02_ncu_full.log 2>&1
echo full rc=$?
