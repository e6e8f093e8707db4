#!/bin/bash
# One parametrised GPU-box script (replaces the numbered one-shot scripts of round 1).
#   gpurun -- 'bash scripts_dev/gpu.sh <mode> [args]'
# modes:
#   probe           which third-party packages the box's image has (scikit-image, tensorflow, ...)
#   verify          whole GPU suite, smoke, both bench arms (what the driver runs at round end)
#   quick [files]   selected GPU tests + one short bench line
#   ab "A=1;B=2"    one short bench line per environment setting
#   launches [env]  ncu launch list (gpu__time_duration) of one 64-patch chunk -> gpurun_out/launches.csv
#   dram            ncu launch list + DRAM bytes of one full 2048^2 step -> gpurun_out/dram_2048.csv
#   ncu <regex> [skip] [count]   ncu --set full of matching kernels in a 64-patch chunk -> gpurun_out/ncu_<tag>.ncu-rep
#   noalloc         tests/test_zz_no_alloc.py alone in a fresh process
#   ncu4            ncu --set full of the stage 3/4 kernels (one C3 image)
#   multi N         bench.py under torchrun on N GPUs, as the driver launches it
#   tests <files>   selected GPU test files only
#   stages          scripts/bench_stages.py (C2/C3 stage table)
mkdir -p gpurun_out
mode=$1; shift
short_bench() {  # $1 = tag, rest = env
  tag=$1; shift
  env "$@" timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/bench_$tag.log 2>&1
  python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_$tag.log').read().strip().splitlines()[-1])
    r=d['roofline']; print('[$tag $*] MP/s', round(d['value'],2), 'ms', round(d['ms_per_step'],3), 'kernel_ms', round(r['kernel_ms_per_step'],3), 'frac', round(r['frac'],3), d['clocks'])
except Exception as e: print('[$tag] bench fail', e)
P
}
case $mode in
probe)
  python - <<'P' 2>&1 | tee gpurun_out/probe.txt
import importlib, subprocess
for m in ("skimage", "tensorflow", "keras", "h5py", "PIL", "cv2", "scipy", "sklearn", "networkx", "numba", "dask", "aicsimageio", "tifffile"):
    try:
        mod = importlib.import_module(m); print(m, "OK", getattr(mod, "__version__", "?"))
    except Exception as e:
        print(m, "ABSENT", type(e).__name__, str(e)[:80])
print(subprocess.run("pip cache dir 2>&1; find / \\( -iname 'scikit_image*' -o -iname 'tensorflow*' \\) -not -path '/proc/*' 2>/dev/null | head -20; ls /opt/wheelhouse 2>/dev/null | head -100; nproc; free -g | head -2", shell=True, capture_output=True, text=True).stdout)
P
  ;;
verify)
  ( time timeout 2400 python -m pytest tests/ -x -q -m gpu ) > gpurun_out/pytest_all.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_all.log
  tail -6 gpurun_out/pytest_all.log
  timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
  timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.log 2>&1; echo "ref exit $?"; tail -1 gpurun_out/bench_ref.log | cut -c1-800
  timeout 900 python bench.py > gpurun_out/bench_default.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench_default.log > gpurun_out/bench_n1.json; cut -c1-3000 gpurun_out/bench_n1.json
  ;;
quick)
  timeout 1200 python -m pytest ${@:-tests/test_gpu_unet.py tests/test_gpu_stp.py} -m gpu -x -q 2>&1 | tail -15 | cut -c1-300
  short_bench q
  ;;
ab)
  IFS=';' read -ra CFG <<< "$1"; k=0
  for c in "${CFG[@]}"; do short_bench ab$k $c; k=$((k+1)); done
  ;;
launches)
  env "$@" TMAT_CHUNK=64 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/ncu_launch.log 2>&1
  echo "ncu exit $?"; python scripts_dev/launch_summary.py gpurun_out/launches.csv | tail -40
  ;;
dram)
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 4000 --csv --log-file gpurun_out/dram_2048.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/ncu_dram.log 2>&1
  echo "ncu dram exit $?"
  ;;
ncu)
  rx=$1; skip=${2:-0}; cnt=${3:-8}; tag=$(echo "$rx" | tr -c 'A-Za-z0-9_\n' '_')
  TMAT_CHUNK=64 timeout 900 ncu --set full --import-source on --clock-control none -k regex:"$rx" --launch-skip $skip --launch-count $cnt -o gpurun_out/ncu_$tag -f python bench.py --size 1024 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/ncu_full.log 2>&1
  echo "ncu full exit $?"
  ;;
noalloc)   # the allocation-free entry point: alone in a fresh process, then again after the whole suite (verify)
  timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -8 | cut -c1-300
  ;;
ncu4)      # ncu --set full of the stage 3/4 kernels of one C3 image (warm caches): the auto_threshold_well sequence and
           # the last filter_branch_seg_mask sequence -> gpurun_out/ncu_stage3.ncu-rep, ncu_stage4.ncu-rep
  RX='regex:thin_|ccl_|erode_disk|root_|gauss|otsu_|median_|footprint_|border_|rescale_u8|minmax64|sk_'
  timeout 600 ncu --cache-control none --set full --import-source on --clock-control none -k "$RX" --launch-skip ${1:-25} --launch-count 5 -o gpurun_out/ncu_stage3 -f python scripts/bench_stages.py --config C3 --no-cpu --iters 2 > gpurun_out/ncu_stage3.log 2>&1
  echo "ncu stage3 exit $?"
  timeout 600 ncu --cache-control none --set full --import-source on --clock-control none -k "$RX" --launch-skip ${2:-103} --launch-count 12 -o gpurun_out/ncu_stage4 -f python scripts/bench_stages.py --config C3 --no-cpu --iters 2 > gpurun_out/ncu_stage4.log 2>&1
  echo "ncu stage4 exit $?"
  ;;
multi)     # bench.py on N GPUs exactly as the driver launches it (gpurun --gpus N): multi N
  N=${1:-2}
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.log 2>&1; echo "bench N=$N exit $?"
  tail -1 gpurun_out/bench_n$N.log > gpurun_out/bench_n$N.json; cut -c1-2500 gpurun_out/bench_n$N.json
  ;;
tests)     # selected GPU test files only (no bench)
  timeout 1500 python -m pytest "$@" -m gpu -x -q 2>&1 | tail -25 | cut -c1-400
  ;;
stages)
  timeout 900 python scripts/bench_stages.py "$@" 2>&1 | tail -40
  ;;
*) echo "unknown mode $mode"; exit 1;;
esac
