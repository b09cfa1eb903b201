#!/bin/bash
cd "$GRAFT_REPO_ROOT"
export B200_SHM_TIMEOUT_S=30
timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu -k "chebyshev_with_a_rank or classical" > gpurun_out/r02i_pytest.log 2>&1; tail -15 gpurun_out/r02i_pytest.log | cut -c1-300
timeout 1500 python profiles/parity_sweep_ranks.py 60 12 > gpurun_out/r02_parity_sweep_ranks_seed12.jsonl 2> gpurun_out/r02_parity_sweep_ranks_seed12.err
echo "ranks sweep rc=$?"; tail -1 gpurun_out/r02_parity_sweep_ranks_seed12.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep_ranks_seed12.jsonl | cut -c1-1500 | head -6
