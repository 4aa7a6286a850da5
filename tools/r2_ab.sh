#!/bin/bash
# A/B of decode schedules on one GPU (run under gpurun): value-only bench legs with different B2_* settings.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
: > gpurun_out/ab.txt
run() { echo "== $*" | tee -a gpurun_out/ab.txt; env "$@" timeout 600 python bench.py --value-only --steps ${STEPS:-120} --warmup 4 2>&1 | tail -1 | tee -a gpurun_out/ab.txt; }
for cfg in ${CFGS:-"B2_MB=0" "B2_MB=32"}; do
  run $cfg
done
