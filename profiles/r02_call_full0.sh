#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 600 python profiles/parity_fullsize.py 256 7 0.25 0 > gpurun_out/r02_parity_256_classical.jsonl 2> gpurun_out/r02_parity_256_classical.err
echo rc=$?; tail -3 gpurun_out/r02_parity_256_classical.jsonl | cut -c1-600; tail -2 gpurun_out/r02_parity_256_classical.err | cut -c1-300
