#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/default_flags.json 2> gpurun_out/default_flags.err; echo "default rc=$?"
timeout 300 python bench.py --steps 3 --warmup 3 --no-extras --cpu-sample-frames 0 > gpurun_out/small_flags.json 2> gpurun_out/small_flags.err; echo "small rc=$?"
python - <<'P'
import json
for f in ("default_flags","small_flags"):
    d=json.load(open(f"gpurun_out/{f}.json")); print(f, d["steps"], d["warmup"], round(d["value"],1), round(d["e2e"]["value"],1), round(d["roofline"]["frac"],4), d["clocks"]["reasons"])
P
