#!/bin/bash
# whole GPU suite on the final commit
cd "$GRAFT_REPO_ROOT"
B200_SHM_TIMEOUT_S=60 timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -8 > gpurun_out/r02w_pytest_gpu.log
cat gpurun_out/r02w_pytest_gpu.log
