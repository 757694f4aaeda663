#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_coarse.py -x -q 2>&1 | tail -3
for rep in 1 2; do
  echo -n "base: "; OPB200_LIB_ALT=tools/probes/base_libopb200.so OPB200_ALGO=0 timeout 120 python tools/sim_debug_time.py 2>&1 | grep "sweep kernel" | sed 's/.*calls)://'
  echo -n "cur : "; OPB200_ALGO=0 OPB200_SIM_DEBUG_COUNT=1 timeout 120 python tools/sim_debug_time.py 2>&1 | grep "sweep kernel\|candidates" | sort | uniq | tail -2 | tr '\n' ' '; echo
done
