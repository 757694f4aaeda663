#!/bin/bash
# full-set ncu capture of the fused encoder kernels (one launch each of the shapes of a 32-frame forward)
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'mlp_fused|attn_merge' --launch-skip 26 -c 8 -o gpurun_out/r02_fused_kernels -f \
  python bench.py --steps 2 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 > gpurun_out/r02_fused_kernels.log 2>&1
echo rc=$?
