#!/bin/bash
# fused MLP kernel (OPB200_MLP_FUSED=1) against the two-launch path at 32 frames per forward, same box, back to back
( OPB200_MLP_FUSED=1 timeout 300 python -m pytest tests/test_gpu_transformer.py tests/test_gpu_models.py -m gpu -x -q ) 2>&1 | tail -3
mkdir -p gpurun_out
OPB200_MLP_FUSED=1 timeout 300 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:'mlp_fused' --launch-skip 12 -c 8 --csv --log-file gpurun_out/mlp_fused_t.csv \
  python bench.py --steps 2 --warmup 3 --no-extras --repeats 1 --cpu-sample-frames 0 > /dev/null 2>&1
grep -E "mlp_fused" gpurun_out/mlp_fused_t.csv | awk -F'","' '{print $(NF-2), $NF}'
for v in 0 1 0 1; do
  echo -n "MLP_FUSED=$v: "; OPB200_MLP_FUSED=$v timeout 600 python bench.py --steps 20 --warmup 5 --no-extras --repeats 3 --cpu-sample-frames 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'single', round(d['single_stream']['value'],1), 'tr256', d['stages_ms_per_frame']['transformer256'], 'tr128', d['stages_ms_per_frame']['transformer128'], 'clk', d['clocks']['sm_mhz'])"
done
