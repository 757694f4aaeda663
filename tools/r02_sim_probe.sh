#!/bin/bash
for rep in 1 2 3; do
for d in 0 9; do
  echo -n "debug $d: "; OPB200_SIM_DEBUG=$d OPB200_ALGO=0 timeout 120 python tools/sim_debug_time.py 2>&1 | grep "sweep kernel" | sed 's/.*calls)://'
done; done
