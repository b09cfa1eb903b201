#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 600 python profiles/run_workloads.py --segregated 256 > gpurun_out/r02_segregated_3rhs.jsonl 2> gpurun_out/r02_segregated_3rhs.err
cat gpurun_out/r02_segregated_3rhs.jsonl; tail -3 gpurun_out/r02_segregated_3rhs.err
timeout 1200 python profiles/parity_sweep.py 150 20261105 > gpurun_out/r02_parity_sweep.jsonl 2> gpurun_out/r02_parity_sweep.err
echo "sweep rc=$?"; tail -2 gpurun_out/r02_parity_sweep.jsonl; grep -c '"ok": true' gpurun_out/r02_parity_sweep.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep.jsonl | head -5; tail -3 gpurun_out/r02_parity_sweep.err
