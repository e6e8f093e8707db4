#!/bin/bash
# first GPU pass of a change: parity tests, chunk-size sweep, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest.log
tail -5 gpurun_out/pytest.log
for ch in 16 32 64 128; do
  TMAT_CHUNK=$ch timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_chunk$ch.log 2>&1
  echo "chunk $ch: $(python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_chunk$ch.log').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_step'], d['roofline']['frac'])
except Exception as e: print('fail', e)
P
)"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_1024.csv python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1
echo "ncu exit $?"
