#!/bin/bash
# launch list of one 64-patch chunk under env settings: SWEEP="A=1;B=2"; outputs gpurun_out/launches_<k>.csv
mkdir -p gpurun_out
IFS=';' read -ra CFG <<< "$SWEEP"
k=0
for c in "${CFG[@]}"; do
  env $c TMAT_CHUNK=64 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_$k.csv python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch_$k.log 2>&1
  echo "[$c] -> launches_$k.csv exit $?"
  k=$((k+1))
done
