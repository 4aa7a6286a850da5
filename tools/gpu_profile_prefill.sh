#!/bin/bash
# ncu launch list (durations only) of the C2 prefill (64 x 512 tokens in 4 chunks of 16 sequences), run under gpurun
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/prof_decode.py --prefill-only > gpurun_out/prof_prefill_plain.log 2>&1 || { tail -5 gpurun_out/prof_prefill_plain.log; exit 1; }
cat gpurun_out/prof_prefill_plain.log
ncu --metrics gpu__time_duration.sum ${EXTRA_METRICS:+,$EXTRA_METRICS} --clock-control none -c 1100 --csv --log-file gpurun_out/${OUT:-r02_prefill_launches}.csv python tools/prof_decode.py --prefill-only > gpurun_out/ncu_prefill.log 2>&1
echo rc=$?; tail -2 gpurun_out/ncu_prefill.log
# MODE=full: one `--set full` capture of a whole layer's heavy kernels (QKV, rope, attention, O, gate/up+SwiGLU, down) of the first chunk
if [ "${MODE:-}" = full ]; then
  ncu --set full --clock-control none --import-source on -k regex:'linear_tc_kernel|linear_mc_kernel|gateup_swiglu_tc_kernel|attn_prefill_tc_kernel|qknorm_rope_kv_prefill16' \
      -s 12 -c 6 -f -o gpurun_out/r02_prefill_full python tools/prof_decode.py --prefill-only > gpurun_out/ncu_prefill_full.log 2>&1
  echo "full rc=$?"; tail -2 gpurun_out/ncu_prefill_full.log
  ncu -i gpurun_out/r02_prefill_full.ncu-rep --page raw --csv > gpurun_out/r02_prefill_full_raw.csv 2>/dev/null
fi
