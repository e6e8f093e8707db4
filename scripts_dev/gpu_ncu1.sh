#!/bin/bash
mkdir -p gpurun_out
TMAT_CHUNK=64 timeout 900 ncu --set full --import-source on --clock-control none -k regex:'conv_halo_kernel|conv_dwpw_kernel' --launch-skip 9 --launch-count 9 -o gpurun_out/ncu_halo_dwpw -f python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"
ls -la gpurun_out/*.ncu-rep
