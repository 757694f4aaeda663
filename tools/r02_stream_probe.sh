#!/bin/bash
# streams per GPU x frames per step at 32 frames per forward, one box, back to back
for cfg in "3 128" "2 128" "4 128" "2 256" "3 256" "3 128"; do
  set -- $cfg
  echo -n "streams $1 (frames/step $2): "; timeout 600 python bench.py --streams $1 --frames-per-step $2 --steps 20 --warmup 5 --no-extras --repeats 3 --cpu-sample-frames 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'roofline', round(d['roofline']['frac'],4), round(d['roofline']['op_level']['frac'],4), 'clk', d['clocks']['sm_mhz'])"
done
