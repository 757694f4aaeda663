#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/stage_bench.py fine > gpurun_out/r02_stage_fine.jsonl 2> gpurun_out/stage_fine.err; echo fine rc=$?
timeout 600 python tools/stage_bench.py coarse --points 5000 8000 30000 > gpurun_out/r02_stage_coarse.jsonl 2> gpurun_out/stage_coarse.err; echo coarse rc=$?
timeout 600 python tools/stage_bench.py coarse --points 5000 30000 --materialize >> gpurun_out/r02_stage_coarse.jsonl 2>> gpurun_out/stage_coarse.err
timeout 600 python tools/stage_bench.py transformer --points 8000 > gpurun_out/r02_stage_transformer.jsonl 2> gpurun_out/stage_tr.err; echo transformer rc=$?
timeout 600 python tools/stage_bench.py conv > gpurun_out/r02_stage_conv.jsonl 2> gpurun_out/stage_conv.err; echo conv rc=$?
timeout 900 python tools/stage_bench.py loftr > gpurun_out/r02_stage_loftr.jsonl 2> gpurun_out/stage_loftr.err; echo loftr rc=$?
tail -n 3 gpurun_out/r02_stage_fine.jsonl | cut -c1-300
