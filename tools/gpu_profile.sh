#!/bin/bash
# ncu evidence (run under gpurun, one GPU): per-launch durations of a steady-state decode step and a
# full-section capture of the dominant kernel.  The plain run must exit 0 first (B200_PROFILING.md).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/prof_plain.log; exit 1; }
tail -1 gpurun_out/prof_plain.log
# one profiler pass per box call: MODE=launches (default) or MODE=full
if [ "${MODE:-launches}" = launches ]; then
ncu --metrics gpu__time_duration.sum --clock-control none -s ${SKIP:-2400} -c ${COUNT:-600} --csv \
    --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
else
ncu --set full --clock-control none --import-source on -k regex:${KERNEL:-attn_decode_kernel} -s ${KSKIP:-100} -c 2 \
    -f -o gpurun_out/prof_${KERNEL:-attn_decode_kernel} $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
fi
ls -la gpurun_out/
