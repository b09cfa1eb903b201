#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 600 python -m pytest tests/test_gpu_app.py -x -q -m gpu -k "generator" 2>&1 | tail -5
timeout 600 python profiles/run_workloads.py --segregated 256 > gpurun_out/r02_segregated_3rhs.jsonl 2> gpurun_out/r02_segregated_3rhs.err
cat gpurun_out/r02_segregated_3rhs.jsonl; tail -3 gpurun_out/r02_segregated_3rhs.err
