#!/bin/bash
OPB200_SIM_DEBUG=0 OPB200_SIM_TRACE=1 timeout 120 python tools/sim_debug_time.py 2>&1 | grep -A31 "^CTA 77"
