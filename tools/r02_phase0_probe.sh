#!/bin/bash
# fine level: MERGE + LN1 as phase 0 of the fused MLP kernel (default) against their own launch (OPB200_MLP_PHASE0=0)
( OPB200_MLP_FUSED=1 timeout 300 python -m pytest tests/test_gpu_transformer.py tests/test_gpu_models.py tests/test_gpu_fine.py -m gpu -x -q ) 2>&1 | tail -3
for v in 0 1 0 1; do
  echo -n "MLP_PHASE0=$v: "; OPB200_MLP_PHASE0=$v timeout 600 python bench.py --steps 20 --warmup 5 --no-extras --repeats 3 --cpu-sample-frames 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'single', round(d['single_stream']['value'],1), 'tr128', d['stages_ms_per_frame']['transformer128'], 'clk', d['clocks']['sm_mhz'])"
done
