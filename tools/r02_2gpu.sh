#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_2gpu.json 2> gpurun_out/r02_bench_2gpu.err
tail -c 400 gpurun_out/r02_bench_2gpu.err
python - <<'P'
import json
d=json.load(open('gpurun_out/r02_bench_2gpu.json'))
print("N=2 value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "strong", d.get("strong_scaling",{}).get("value"), d.get("strong_scaling",{}).get("checksum"), "loftr", d.get("secondary",{}).get("loftr_pairs",{}).get("value"))
print(d["diag"])
P
