#!/bin/bash
# sim_fast_kernel: pairs of tiles on 128 x 256 MMAs (OPB200_SIM_PAIR=1) against the 128-wide schedule, same box
( OPB200_SIM_PAIR=1 timeout 300 python -m pytest tests/test_gpu_coarse.py -m gpu -x -q ) 2>&1 | tail -3
for n in 8 32; do for v in 0 1 0 1; do
  echo -n "N=$n PAIR=$v: "; OPB200_N=$n OPB200_SIM_PAIR=$v timeout 120 python tools/sim_debug_time.py 2>&1 | grep "sweep kernel" | sed 's/.*sweep kernel us/us/'
done; done
