#!/bin/bash
# guarded check of a risky kernel change: dwpw tests under a short timeout, then the unet tests, then one bench line
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_unet.py -m gpu -x -q -k "fused_separable or fused_maxpool" 2>&1 | tail -15 | cut -c1-300
echo "dwpw exit ${PIPESTATUS[0]}"
timeout 180 python -m pytest tests/test_gpu_unet.py tests/test_gpu_stp.py -m gpu -x -q 2>&1 | tail -4 | cut -c1-300
timeout 120 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_g.log 2>&1
python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_g.log').read().strip().splitlines()[-1]); print('bench', round(d['value'],2), round(d['ms_per_step'],3), round(d['roofline']['kernel_ms_per_step'],3), d['clocks'])
except Exception as e: print('bench fail', e)
P
