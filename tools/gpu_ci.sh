#!/bin/bash
# GPU-box check list (run under gpurun): every stage is bounded by `timeout` so a hung kernel
# cannot hold the box, and logs land in gpurun_out/ so they come back.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
PY=python
run() { # name timeout cmd...
  local name=$1 t=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/ci_summary.txt
  timeout -k 10 "$t" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?
  echo "rc=$rc" | tee -a gpurun_out/ci_summary.txt
  tail -n 15 "gpurun_out/$name.log" | tee -a gpurun_out/ci_summary.txt
  return $rc
}
: > gpurun_out/ci_summary.txt
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.sm --format=csv | tee -a gpurun_out/ci_summary.txt
STAGES=${STAGES:-"elem tc attn engine_simt engine smoke bench"}
for st in $STAGES; do
case $st in
  elem)   run k_elem 600 $PY -m pytest tests/test_kernels_gpu.py -q --tb=short -p no:cacheprovider --timeout=240 --timeout-method=thread -k "rmsnorm or qknorm or silu or (linear_bf16 and 0-)" ;;
  tc)     run k_tc 900 $PY -m pytest tests/test_kernels_gpu.py -q --tb=short -p no:cacheprovider --timeout=240 --timeout-method=thread -k "linear or lmhead" ;;
  attn)   run k_attn 600 $PY -m pytest tests/test_kernels_gpu.py -q --tb=short -p no:cacheprovider --timeout=240 --timeout-method=thread -k "attn" ;;
  engine_simt) B2_GEMM_IMPL=0 run e_simt 900 $PY -m pytest tests/test_engine_gpu.py -q --tb=short -p no:cacheprovider --timeout=400 --timeout-method=thread -k "tiny" ;;
  engine) run e_full 1500 $PY -m pytest tests/test_engine_gpu.py -q --tb=short -p no:cacheprovider --timeout=600 --timeout-method=thread -s ;;
  smoke)  run smoke 300 $PY -c "import __graft_entry__ as g; g.smoke()" ;;
  bench)  run bench 900 $PY bench.py ${BENCH_ARGS:-} ;;
  all)    run all_gpu 2400 $PY -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout=600 --timeout-method=thread ;;
esac
done
echo "=== done" | tee -a gpurun_out/ci_summary.txt
