#!/bin/bash
cd "$GRAFT_REPO_ROOT"
FAIL=""
for n in 24 48 96 160 256; do
  timeout 600 python profiles/dbg_poisson.py $n > gpurun_out/dbg_p$n.log 2>&1; rc=$?
  tail -2 gpurun_out/dbg_p$n.log | cut -c1-300
  if [ $rc -ne 0 ] && [ -z "$FAIL" ]; then FAIL=$n; fi
done
echo "first failing size: $FAIL"
if [ -n "$FAIL" ]; then
  timeout 1500 compute-sanitizer --tool memcheck --print-limit 20 python profiles/dbg_poisson.py $FAIL > gpurun_out/dbg_memcheck.log 2>&1
  grep -v "^$" gpurun_out/dbg_memcheck.log | head -60 | cut -c1-400
fi
