#!/bin/bash
# ncu --set full capture of selected launches of one bench step (run under gpurun):
#   bash profiles/prof_kernel.sh <tag> <kernel-regex> <skip> <count> [extra bench args]
TAG=$1; RE=$2; SKIP=${3:-0}; CNT=${4:-1}; shift 4
mkdir -p gpurun_out
CMD="python bench.py --size 256 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline $*"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$RE" -s $SKIP -c $CNT -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
tail -3 gpurun_out/ncu_full_$TAG.log
