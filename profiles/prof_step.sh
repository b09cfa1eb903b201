#!/bin/bash
# ncu recipe of /opt/skills/guides/B200_PROFILING.md for one bench step at the headline size (run under gpurun):
#   bash profiles/prof_step.sh <tag> [kernel-regex] [skip] [count]
# writes gpurun_out/launches_<tag>.csv (every launch with its device time) and, when a regex is given,
# gpurun_out/prof_<tag>.ncu-rep (--set full of the matching launches).
TAG=${1:-r01}
RE=${2:-}
SKIP=${3:-0}
CNT=${4:-3}
mkdir -p gpurun_out
CMD="python bench.py --size 256 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --parity-n 0"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
if [ -n "$RE" ]; then
  ncu --set full --clock-control none --import-source on -k "regex:$RE" -s $SKIP -c $CNT -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
fi
tail -2 gpurun_out/plain_$TAG.log | cut -c1-600
tail -3 gpurun_out/ncu_list_$TAG.log gpurun_out/ncu_full_$TAG.log 2>/dev/null
ls -la gpurun_out
