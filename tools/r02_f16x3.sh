#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 --backbone f16x3 > gpurun_out/r02_bench_f16x3.json 2> gpurun_out/r02_bench_f16x3.err
python - <<'P'
import json
d=json.load(open('gpurun_out/r02_bench_f16x3.json'))
print("value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "roofline", round(d["roofline"]["frac"],4))
print(d["stages_ms_per_frame"]); print("parity", d.get("parity"))
P
