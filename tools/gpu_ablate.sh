#!/bin/bash
# Timing ablation (run under gpurun): leave one kernel class out of the captured decode graph and
# read the change in ms/step -- the class's share of the critical path with PDL overlap in place.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
: > gpurun_out/ablate.txt
# classes: 0 embed 1 norm 2 gemm_qkv 3 rope 4 attn 5 combine 6 gemm_o 7 gemm_gu 8 act 9 gemm_down 10 lmhead 11 misc
for m in ${MASKS:-0 2 4 8 16 32 64 128 256 512 1024 708 1006}; do
  B2_SKIP_MASK=$m python bench.py --value-only --steps ${STEPS:-120} --warmup 4 2>&1 | tail -1 | tee -a gpurun_out/ablate.txt
done
