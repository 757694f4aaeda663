#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_transformer.py tests/test_gpu_models.py tests/test_gpu_bench_config.py -x -q 2>&1 | tail -3
for v in 0 1 0 1; do
  echo -n "MLP_FUSED=$v: "; OPB200_MLP_FUSED=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'transformer256', d['stages_ms_per_frame']['transformer256'], 'single', round(d['single_stream']['value'],1))"
done
