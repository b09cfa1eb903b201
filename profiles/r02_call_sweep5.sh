#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 1200 python profiles/parity_sweep_unstructured.py 60 3 > gpurun_out/r02_parity_sweep_unstructured.jsonl 2> gpurun_out/r02_parity_sweep_unstructured.err
echo "unstructured sweep rc=$?"; tail -1 gpurun_out/r02_parity_sweep_unstructured.jsonl; grep '"ok": false' gpurun_out/r02_parity_sweep_unstructured.jsonl | cut -c1-900 | head -8; tail -4 gpurun_out/r02_parity_sweep_unstructured.err | cut -c1-400
