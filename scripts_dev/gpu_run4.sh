#!/bin/bash
mkdir -p gpurun_out
timeout 120 python scripts_dev/halo_probe.py 1 > gpurun_out/halo_probe_1.log 2>&1; tail -1 gpurun_out/halo_probe_1.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest.log
tail -5 gpurun_out/pytest.log
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_$name.log 2>&1
  echo "$name: $(python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_$name.log').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_step'], d['roofline']['frac'])
except Exception as e: print('fail', e)
P
)"
}
run default
run chunk128 TMAT_CHUNK=128
run chunk512 TMAT_CHUNK=512
TMAT_CHUNK=64 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_1024.csv python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1
echo "ncu exit $?"
