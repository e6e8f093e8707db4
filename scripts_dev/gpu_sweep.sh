#!/bin/bash
# bench.py under a list of env settings: SWEEP="A=1;A=2 B=3;..."
mkdir -p gpurun_out
IFS=';' read -ra CFG <<< "$SWEEP"
for c in "${CFG[@]}"; do
  env $c timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_sw.log 2>&1
  python - "$c" <<P
import json,sys
try:
    d=json.loads(open('gpurun_out/bench_sw.log').read().strip().splitlines()[-1]); print('bench[%s]'%sys.argv[1], round(d['value'],2), round(d['ms_per_step'],2), round(d['roofline']['kernel_ms_per_step'],2), d['clocks']['sm_mhz'])
except Exception as e: print('fail', sys.argv[1], e)
P
done
