#!/bin/bash
# ncu --set full of the classical row kernel on levels 0 and 1 of lap7_256 (the plain command exited 0 in the previous call)
cd "$GRAFT_REPO_ROOT"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_classical -c 4 -o gpurun_out/r02c0_classical python profiles/dbg_classical.py 256 0 > gpurun_out/r02c0_classical_ncu.log 2>&1
tail -3 gpurun_out/r02c0_classical_ncu.log
ls -la gpurun_out/*.ncu-rep
