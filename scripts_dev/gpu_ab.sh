#!/bin/bash
# A/B of env settings: AB="A=1;B=2" -> dwpw/unet tests + one bench line each
mkdir -p gpurun_out
IFS=';' read -ra CFG <<< "$AB"
k=0
for c in "${CFG[@]}"; do
  env $c timeout 600 python -m pytest tests/test_gpu_unet.py -m gpu -x -q 2>&1 | tail -1
  env $c timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_ab_$k.log 2>&1
  python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_ab_$k.log').read().strip().splitlines()[-1]); print('[$c] bench', round(d['value'],2), round(d['ms_per_step'],3), round(d['roofline']['kernel_ms_per_step'],3), d['clocks'])
except Exception as e: print('[$c] fail', e)
P
  k=$((k+1))
done
