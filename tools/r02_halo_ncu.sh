#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none -k regex:'conv_kernel' --launch-skip 42 -c 21 -o gpurun_out/r02_conv_halo -f \
  python bench.py --steps 2 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 > gpurun_out/r02_conv_halo.log 2>&1
echo rc=$?
