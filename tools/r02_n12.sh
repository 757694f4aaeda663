#!/bin/bash
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) 2>&1 | tail -2
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/n1.json 2> gpurun_out/n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/n2.json 2> gpurun_out/n2.err
python - <<'P'
import json
for f in ("n1","n2"):
    d=json.load(open(f"gpurun_out/{f}.json"))
    print(f, "value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "roofline", round(d["roofline"]["frac"],4), "strong", round(d["strong_scaling"]["value"],1), d["strong_scaling"]["checksum"], d["strong_scaling"]["matches"], "parity", d.get("parity",{}).get("mconf_max_abs_err"), "cg", (d.get("contract_grade") or {}).get("parity",{}).get("green"))
P
