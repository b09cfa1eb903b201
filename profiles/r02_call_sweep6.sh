#!/bin/bash
cd "$GRAFT_REPO_ROOT"
export B200_SHM_TIMEOUT_S=30
timeout 1500 python profiles/parity_sweep_unstructured.py --ranks 36 4 > gpurun_out/r02_parity_sweep_unstructured_ranks.jsonl 2> gpurun_out/r02_parity_sweep_unstructured_ranks.err
echo "unstructured ranks sweep rc=$?"; tail -1 gpurun_out/r02_parity_sweep_unstructured_ranks.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep_unstructured_ranks.jsonl | cut -c1-1800 | head -6; tail -4 gpurun_out/r02_parity_sweep_unstructured_ranks.err | cut -c1-400
