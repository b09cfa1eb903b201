#!/bin/bash
cd "$GRAFT_REPO_ROOT"
export B200_SHM_TIMEOUT_S=30
timeout 900 python profiles/parity_sweep_ranks.py 40 13 > gpurun_out/r02_parity_sweep_ranks_seed13.jsonl 2> gpurun_out/r02_parity_sweep_ranks_seed13.err
echo "ranks sweep rc=$?"; tail -1 gpurun_out/r02_parity_sweep_ranks_seed13.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep_ranks_seed13.jsonl | cut -c1-1500 | head -6
