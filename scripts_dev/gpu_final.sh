#!/bin/bash
# round-end evidence in one call: whole GPU suite, smoke, both bench arms, launch list + DRAM bytes of one step,
# full-set capture of the fused SeparableConv2D kernels
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/ -x -q -m gpu ) > gpurun_out/pytest_all.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_all.log
tail -6 gpurun_out/pytest_all.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench_default.log > gpurun_out/bench_n1.json; cut -c1-400 gpurun_out/bench_n1.json
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.log 2>&1; echo "ref exit $?"; tail -1 gpurun_out/bench_ref.log | cut -c1-300
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 3000 --csv --log-file gpurun_out/dram_2048.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_dram.log 2>&1
echo "ncu dram exit $?"
TMAT_CHUNK=64 timeout 600 ncu --set full --import-source on --clock-control none -k regex:'conv_dwpw_kernel' --launch-skip 6 --launch-count 6 -o gpurun_out/ncu_dwpw6 -f python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
