#!/bin/bash
cd "$GRAFT_REPO_ROOT"
export B200_SHM_TIMEOUT_S=45
timeout 600 python profiles/dbg_ranks_case.py '[2,2,2]' '[8,5,5]' 27 0.5 0 4 0.0 > gpurun_out/dbg_case15.log 2>&1
grep -v "^\s*File\|^    " gpurun_out/dbg_case15.log | tail -12
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "host_rows_cross" > gpurun_out/r02h_pytest_hostrows.log 2>&1; tail -12 gpurun_out/r02h_pytest_hostrows.log
timeout 1500 python profiles/parity_sweep_ranks.py 40 20261106 > gpurun_out/r02_parity_sweep_ranks.jsonl 2> gpurun_out/r02_parity_sweep_ranks.err
echo "ranks sweep rc=$?"; tail -1 gpurun_out/r02_parity_sweep_ranks.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep_ranks.jsonl | cut -c1-400 | head -5
