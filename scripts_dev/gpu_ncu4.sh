#!/bin/bash
# full ncu capture (with source) of stem + first maxpool of one chunk
mkdir -p gpurun_out
TMAT_CHUNK=64 timeout 900 ncu --set full --import-source on --clock-control none --cache-control none -k regex:'stem_kernel|maxpool_kernel' --launch-skip 4 --launch-count 2 -o gpurun_out/ncu_stem4 -f python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"
