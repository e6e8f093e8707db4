#!/bin/bash
# sanity tests, then DRAM bytes + duration of every launch of one 2048x2048 step (for roofline.traffic)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_unet.py tests/test_gpu_stp.py -m gpu -x -q > gpurun_out/pytest.log 2>&1; tail -2 gpurun_out/pytest.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_q.log 2>&1; tail -1 gpurun_out/bench_q.log | cut -c1-200
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 3000 --csv --log-file gpurun_out/dram_2048.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_dram.log 2>&1
echo "ncu exit $?"
