#!/bin/bash
# other BASELINE configurations and the sibling Krylov methods on the final sources (one B200)
cd "$GRAFT_REPO_ROOT"
timeout 420 python profiles/run_workloads.py > gpurun_out/r02w_workloads.jsonl 2> gpurun_out/r02w_workloads.err
echo rc=$?; cut -c1-260 gpurun_out/r02w_workloads.jsonl; tail -2 gpurun_out/r02w_workloads.err | cut -c1-300
