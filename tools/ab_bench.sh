#!/bin/bash
# A/B of bench.py under different environments on ONE box (boxes differ by +-1.5 %):  tools/ab_bench.sh "A=1" "A=0" ...
cd "$(dirname "$0")/.."
for cfg in "$@"; do
  env $cfg timeout 300 python bench.py --steps ${STEPS:-20} --warmup 3 2>/dev/null | tail -1 | \
    python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$cfg', 'tok/s %.0f' % d['value'], 'ms/step %.4f' % d['ms_per_step'], 'prefill_ms %.2f' % (1e3 * d['prefill_s']), 'launches', d['gpu_launches'])"
done
