#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/n4.json 2> gpurun_out/n4.err
python - <<'P'
import json
d=json.load(open("gpurun_out/n4.json"))
print("n4 value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "roofline", round(d["roofline"]["frac"],4), "strong", round(d["strong_scaling"]["value"],1), d["strong_scaling"]["checksum"], d["strong_scaling"]["matches"], "loftr", round(d["secondary"]["loftr_pairs"]["value"],1))
P
