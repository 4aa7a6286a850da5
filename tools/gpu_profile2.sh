#!/bin/bash
# ncu evidence for one decode step at the bench shape (run under gpurun, one GPU).  The plain run must exit 0 first.
#   MODE=launches : per-launch durations of two steady-state steps          -> gpurun_out/r02_launches.csv
#   MODE=full KERNEL=<regex> : --set full capture of that kernel's launches -> gpurun_out/r02_<name>.ncu-rep (+ raw csv)
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python tools/prof_decode.py --ctx ${CTX:-640} --steps 2"
$CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/prof_plain.log; exit 1; }
cat gpurun_out/prof_plain.log
SKIP=$(grep "before the profiled" gpurun_out/prof_plain.log | sed -E 's/.*: ([0-9]+)/\1/')
# the graph-replayed steps appear to ncu as their kernel nodes; eager launches (prefill, graph warm-up) come first
echo "skipping $SKIP launches"
if [ "${MODE:-launches}" = launches ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c ${COUNT:-520} --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
  echo "launch list rc=$?"; tail -3 gpurun_out/ncu_launches.log
else
  NAME=${NAME:-$KERNEL}
  ncu --set full --clock-control none --import-source on -k regex:$KERNEL -s ${KSKIP:-0} -c ${KCOUNT:-4} -f -o gpurun_out/r02_$NAME $CMD > gpurun_out/ncu_full_$NAME.log 2>&1
  echo "full capture rc=$?"; tail -3 gpurun_out/ncu_full_$NAME.log
  ncu -i gpurun_out/r02_$NAME.ncu-rep --page raw --csv > gpurun_out/r02_${NAME}_raw.csv 2>/dev/null
fi
ls -la gpurun_out/ | tail -12
