#!/bin/bash
# ncu launch list (durations only) of the C2 prefill (64 x 512 tokens in 4 chunks of 16 sequences), run under gpurun
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/prof_decode.py --prefill-only > gpurun_out/prof_prefill_plain.log 2>&1 || { tail -5 gpurun_out/prof_prefill_plain.log; exit 1; }
cat gpurun_out/prof_prefill_plain.log
ncu --metrics gpu__time_duration.sum ${EXTRA_METRICS:+,$EXTRA_METRICS} --clock-control none -c 1100 --csv --log-file gpurun_out/${OUT:-r02_prefill_launches}.csv python tools/prof_decode.py --prefill-only > gpurun_out/ncu_prefill.log 2>&1
echo rc=$?; tail -2 gpurun_out/ncu_prefill.log
