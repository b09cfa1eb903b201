#!/bin/bash
cd "$GRAFT_REPO_ROOT"
export B200_SHM_TIMEOUT_S=45
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "host_rows_cross or ij_assembly" > gpurun_out/r02h_pytest_hostrows.log 2>&1; tail -12 gpurun_out/r02h_pytest_hostrows.log
