#!/bin/bash
mkdir -p gpurun_out
TMAT_CHUNK=64 timeout 900 ncu --set full --import-source on --clock-control none --cache-control none -k regex:'conv_halo_kernel' --launch-skip 3 --launch-count 3 -o gpurun_out/ncu_halo2 -f python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"
