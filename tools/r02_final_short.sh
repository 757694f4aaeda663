#!/bin/bash
# tests, smoke and the N=1 bench line (no ncu passes)
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests -m gpu -x -q ) 2>&1 | tail -2 | tee gpurun_out/final_tests.txt
( timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke OK')" ) 2>&1 | tail -1 | tee -a gpurun_out/final_tests.txt
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/n1.json 2> gpurun_out/n1.err; echo bench rc=$?
python tools/show_bench.py < gpurun_out/n1.json 2>/dev/null | head -3
