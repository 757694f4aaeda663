#!/bin/bash
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_coarse.py tests/test_gpu_bench_config.py -m gpu -x -q ) 2>&1 | tail -2
for n in 8 32; do
OPB200_N=$n timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'merge_fast|pack_f16|colmean|sim_fast|cand_|compact' --launch-skip 32 -c 16 --csv --log-file gpurun_out/merge_$n.csv python tools/sim_debug_time.py > /dev/null 2>&1
true
done
