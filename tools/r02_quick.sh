#!/bin/bash
# quick GPU check: full parity suite, smoke, one bench line
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) 2>&1 | tail -3
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" ) 2>&1 | tail -3
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/quick_bench.json 2> gpurun_out/quick_bench.err
python - <<'P'
import json
d=json.load(open('gpurun_out/quick_bench.json'))
print("value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "roofline", round(d["roofline"]["frac"],4), "kernel ms", round(d["roofline"]["launch"]["avg_ms_per_launch"],5), "op", round(d["roofline"]["op_level"].get("frac",0),4))
print(d["stages_ms_per_frame"]); print("parity", d.get("parity"))
P
for d in 0 1; do timeout 120 python tools/clock_probe.py $d 2>&1 | tail -1; done
