#!/bin/bash
# unstructured parity tests of the suite, then the ncu launch list of one bench step on the final sources
cd "$GRAFT_REPO_ROOT"
export B200_SHM_TIMEOUT_S=30
timeout 900 python -m pytest tests/test_gpu_unstructured.py -x -q -m gpu > gpurun_out/r02w_pytest_unstructured.log 2>&1; tail -6 gpurun_out/r02w_pytest_unstructured.log | cut -c1-300
bash profiles/prof_step.sh r02w > gpurun_out/r02w_prof_step.log 2>&1
ls -la gpurun_out/launches_r02w.csv
