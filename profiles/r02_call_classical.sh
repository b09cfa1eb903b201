#!/bin/bash
# classical interpolation (interp_type 0) and halo_map: reference -- parity tests on 1 and N ranks, the sample input through both
# hosts, timing of the classical setup at 256^3
cd "$GRAFT_REPO_ROOT"
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multirank.py tests/test_gpu_app.py -x -q -m gpu -k "classical or sample or halo_map" 2>&1 | tail -40 > gpurun_out/r02c0_pytest_classical.log
cat gpurun_out/r02c0_pytest_classical.log
timeout 600 python profiles/dbg_classical.py 256 0 > gpurun_out/r02c0_classical_256.log 2>&1
tail -5 gpurun_out/r02c0_classical_256.log
