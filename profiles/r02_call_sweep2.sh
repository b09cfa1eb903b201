#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 1500 python profiles/parity_sweep_ranks.py 40 20261106 > gpurun_out/r02_parity_sweep_ranks.jsonl 2> gpurun_out/r02_parity_sweep_ranks.err
echo "ranks sweep rc=$?"; tail -1 gpurun_out/r02_parity_sweep_ranks.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep_ranks.jsonl | head -5; tail -3 gpurun_out/r02_parity_sweep_ranks.err
timeout 1200 python profiles/parity_sweep.py 300 7 > gpurun_out/r02_parity_sweep_seed7.jsonl 2> gpurun_out/r02_parity_sweep_seed7.err
echo "sweep rc=$?"; tail -1 gpurun_out/r02_parity_sweep_seed7.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep_seed7.jsonl | head -5
