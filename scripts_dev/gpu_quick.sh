#!/bin/bash
# quick check of a kernel change: conv self-tests + unet tests, one bench line, launch list of a 64-patch chunk
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_unet.py tests/test_gpu_stp.py -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest.log
tail -3 gpurun_out/pytest.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_q.log 2>&1
python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_q.log').read().strip().splitlines()[-1]); print('bench', d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_step'], d['roofline']['frac'], d['clocks'])
except Exception as e: print('fail', e)
P
TMAT_CHUNK=64 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_1024.csv python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1
echo "ncu exit $?"
if [ -n "$QUICK_EXTRA_ENV" ]; then
  env $QUICK_EXTRA_ENV timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_q2.log 2>&1
  python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_q2.log').read().strip().splitlines()[-1]); print('bench[extra env]', d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_step'])
except Exception as e: print('fail', e)
P
fi
