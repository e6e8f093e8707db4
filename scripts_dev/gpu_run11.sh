#!/bin/bash
# fused maxpool epilogue: bit-identity self-tests, UNet/stp tests, bench with and without the fusion
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_unet.py -m gpu -x -q -k "maxpool_epilogue" > gpurun_out/pytest_pool.log 2>&1; echo "pool exit $?"
tail -15 gpurun_out/pytest_pool.log
timeout 900 python -m pytest tests/test_gpu_unet.py tests/test_gpu_stp.py -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/pytest.log
for env in "X=1" "TMAT_NO_FUSE_POOL=1"; do
env $env timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_q.log 2>&1
python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_q.log').read().strip().splitlines()[-1]); print('bench[$env]', d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_step'], d['roofline']['frac'], d['clocks'], d['image_metrics'])
except Exception as e: print('fail', e); print(open('gpurun_out/bench_q.log').read()[-1500:])
P
done
