#!/bin/bash
# one-lane step timelines under different B2_* settings (run under gpurun): summary table + one layer
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
i=0
for cfg in "$@"; do
  i=$((i+1))
  echo "=== $cfg"
  env $cfg timeout 300 python tools/step_timeline.py --out gpurun_out/tl1_$i.json > gpurun_out/tl1_$i.txt 2>&1
  head -${LINES:-28} gpurun_out/tl1_$i.txt
done
