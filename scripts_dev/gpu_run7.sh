#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n2.log 2>&1; tail -1 gpurun_out/bench_n2.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/bench_c5_rowshard.py --size 2048 --steps 2 --check > gpurun_out/c5_check_n2.log 2>&1; tail -1 gpurun_out/c5_check_n2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 scripts/bench_c5_rowshard.py --size 8192 --steps 2 > gpurun_out/c5_n2.log 2>&1; tail -1 gpurun_out/c5_n2.log
timeout 900 python scripts/bench_c5_rowshard.py --size 8192 --steps 2 > gpurun_out/c5_n1.log 2>&1; tail -1 gpurun_out/c5_n1.log
