#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stp.py tests/test_gpu_unet.py -m gpu -x -q > gpurun_out/pytest.log 2>&1; tail -3 gpurun_out/pytest.log
timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/bench_e2e.log 2>&1
python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_e2e.log').read().strip().splitlines()[-1]); print('bench', round(d['value'],2), round(d['ms_per_step'],2), 'e2e', d['e2e'], d['roofline']['traffic'], d['clocks'])
except Exception as e: print('fail', e); print(open('gpurun_out/bench_e2e.log').read()[-1500:])
P
