#!/bin/bash
# where the fused MLP kernel starts to pay: frames per forward x forced off / on
for b in 8 16; do for v in 0 1; do
  echo -n "batch $b MLP_FUSED=$v: "; OPB200_MLP_FUSED=$v timeout 600 python bench.py --batch $b --steps 20 --warmup 5 --no-extras --repeats 3 --cpu-sample-frames 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'single', round(d['single_stream']['value'],1), 'tr256', d['stages_ms_per_frame']['transformer256'], 'clk', d['clocks']['sm_mhz'])"
done; done
