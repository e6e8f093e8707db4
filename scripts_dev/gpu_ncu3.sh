#!/bin/bash
# full ncu capture (with source) of the fused depthwise+pointwise kernels of one chunk: d0.sep1, d0.sep2, d1.sep1
mkdir -p gpurun_out
TMAT_CHUNK=64 timeout 900 ncu --set full --import-source on --clock-control none -k regex:'conv_dwpw_kernel' --launch-skip 6 --launch-count 3 -o gpurun_out/ncu_dwpw3 -f python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"
