#!/bin/bash
for cfg in "8 128" "16 128" "32 128" "24 120" "16 128" "32 128"; do
  set -- $cfg
  echo -n "batch $1 (frames/step $2): "; timeout 600 python bench.py --batch $1 --frames-per-step $2 --steps 20 --warmup 5 --no-extras --repeats 3 --cpu-sample-frames 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'single', round(d['single_stream']['value'],1), 'roofline', round(d['roofline']['frac'],4), round(d['roofline']['op_level']['frac'],4), 'tr256', d['stages_ms_per_frame']['transformer256'])"
done
nvidia-smi --query-gpu=memory.used --format=csv
