#!/bin/bash
# halo boxes of the 3x3 stride-1 convolutions (default) against one box per tap (OPB200_CONV_HALO=0), same box
( timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_fine.py tests/test_gpu_models.py tests/test_gpu_pose.py -m gpu -x -q ) 2>&1 | tail -3
for v in 0 1 0 1; do
  echo -n "CONV_HALO=$v: "; OPB200_CONV_HALO=$v timeout 600 python bench.py --steps 20 --warmup 5 --no-extras --repeats 3 --cpu-sample-frames 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'single', round(d['single_stream']['value'],1), 'conv', d['stages_ms_per_frame']['backbone_conv'], 'clk', d['clocks']['sm_mhz'])"
done
