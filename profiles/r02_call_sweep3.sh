#!/bin/bash
cd "$GRAFT_REPO_ROOT"
export B200_SHM_TIMEOUT_S=30
timeout 1200 python profiles/parity_sweep.py 200 11 > gpurun_out/r02_parity_sweep_seed11.jsonl 2> gpurun_out/r02_parity_sweep_seed11.err
echo "sweep rc=$?"; tail -1 gpurun_out/r02_parity_sweep_seed11.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep_seed11.jsonl | cut -c1-500 | head -8; tail -3 gpurun_out/r02_parity_sweep_seed11.err | cut -c1-300
timeout 1500 python profiles/parity_sweep_ranks.py 40 11 > gpurun_out/r02_parity_sweep_ranks_seed11.jsonl 2> gpurun_out/r02_parity_sweep_ranks_seed11.err
echo "ranks sweep rc=$?"; tail -1 gpurun_out/r02_parity_sweep_ranks_seed11.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep_ranks_seed11.jsonl | cut -c1-1200 | head -6
