#!/bin/bash
# ncu launch list of one bench command + full-set capture of the CoarseMatching kernels (final round-2 code)
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err; echo plain rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_bench_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 > gpurun_out/r02_ncu_launch.log 2>&1
echo launches rc=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'sim_fast_kernel|merge_fast|pack_f16|colmean|cand_|compact' --launch-skip 48 -c 9 -o gpurun_out/r02_coarse_full -f \
  python bench.py --steps 2 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 > gpurun_out/r02_ncu_full.log 2>&1
echo full rc=$?
