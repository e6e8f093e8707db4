#!/bin/bash
# two-GPU refresh: bench.py at N=2 (image-sharded), C5 row strips (bit-identity check at 2048, 8192 timing)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 400 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.log 2>&1; echo "bench n2 exit $?"; tail -1 gpurun_out/bench_n2.log > gpurun_out/bench_n2.json; cut -c1-300 gpurun_out/bench_n2.json
: > gpurun_out/c5.jsonl
timeout 300 $TR scripts/bench_c5_rowshard.py --size 2048 --check 2>&1 | tail -1 >> gpurun_out/c5.jsonl
timeout 300 python scripts/bench_c5_rowshard.py --size 8192 2>&1 | tail -1 >> gpurun_out/c5.jsonl
timeout 300 $TR scripts/bench_c5_rowshard.py --size 8192 2>&1 | tail -1 >> gpurun_out/c5.jsonl
cut -c1-330 gpurun_out/c5.jsonl
