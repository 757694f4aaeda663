#!/bin/bash
# full-set capture of the fine-level kernels (128-wide GEMMs, attn_small, gather, match) of one forward
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'gemm_kernel<.int.128' --launch-skip 24 -c 8 -o gpurun_out/r02_fine_gemm_full -f \
  python bench.py --steps 2 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 > gpurun_out/r02_ncu_fine.log 2>&1
echo rc=$?
