#!/bin/bash
# usage: [GPUS=N] profiles/gpurun_retry.sh <timeout_s> <logfile> <command...>; retries while gpurun answers "no box / slot free" (exit 3)
T=$1; LOG=$2; shift 2
G=""
if [ -n "$GPUS" ]; then G="--gpus $GPUS"; fi
for k in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $G --timeout "$T" -- "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
