#!/bin/bash
mkdir -p gpurun_out
for d in 12 8 4 16 2 0; do
echo "== TMAT_POOL_DEBUG=$d"
TMAT_POOL_DEBUG=$d timeout 300 python -m pytest tests/test_gpu_unet.py -m gpu -x -q -k "maxpool_epilogue and case4" 2>&1 | grep -i "error\|assert \|passed\|failed" | cut -c1-250 | head -4
done
