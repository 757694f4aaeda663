#!/bin/bash
# threshold change: the 8-frame bench-configuration test now runs the fused encoder kernels (760 tiles >= 4 x 148)
( timeout 600 python -m pytest tests/test_gpu_bench_config.py tests/test_gpu_transformer.py tests/test_gpu_models.py tests/test_gpu_pose.py -m gpu -x -q ) 2>&1 | tail -2
for b in 8; do
  echo -n "batch $b: "; timeout 300 python bench.py --batch $b --steps 10 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'tr256', d['stages_ms_per_frame']['transformer256'])"
done
