#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 900 python profiles/parity_sweep_krylov.py 100 5 > gpurun_out/r02_parity_sweep_krylov.jsonl 2> gpurun_out/r02_parity_sweep_krylov.err
echo "krylov sweep rc=$?"; tail -1 gpurun_out/r02_parity_sweep_krylov.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep_krylov.jsonl | cut -c1-700 | head -8; tail -3 gpurun_out/r02_parity_sweep_krylov.err | cut -c1-300
