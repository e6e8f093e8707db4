#!/bin/bash
# end-of-round check: all GPU tests, bench line (with e2e + CPU arm), stage benches C2/C3/C4
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest.log
tail -3 gpurun_out/pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_full.log 2>&1
tail -c 3000 gpurun_out/bench_full.log
timeout 900 python scripts/bench_stages.py > gpurun_out/stages.jsonl 2> gpurun_out/stages.err
echo "stages exit $?"; cat gpurun_out/stages.jsonl | cut -c1-420; tail -5 gpurun_out/stages.err
