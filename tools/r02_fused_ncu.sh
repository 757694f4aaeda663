#!/bin/bash
mkdir -p gpurun_out
OPB200_MLP_FUSED=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'mlp_fused' --launch-skip 12 -c 3 -o gpurun_out/r02_mlp_fused -f \
  python bench.py --steps 2 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 > gpurun_out/r02_mlp_fused.log 2>&1
echo rc=$?
