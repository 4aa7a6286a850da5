#!/bin/bash
# step timelines under different B2_* settings (run under gpurun); prints the both-lanes window of each
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
i=0
for cfg in "$@"; do
  i=$((i+1))
  echo "=== $cfg"
  env $cfg timeout 300 python tools/step_timeline.py --out gpurun_out/tl_$i.json > gpurun_out/tl_$i.txt 2>&1
  head -1 gpurun_out/tl_$i.txt
  grep -A${LINES:-24} "both lanes" gpurun_out/tl_$i.txt
  tail -2 gpurun_out/tl_$i.txt
done
