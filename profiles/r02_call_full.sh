#!/bin/bash
# round-2: BASELINE.json configs[3] / [4] at their stated sizes, generated on the device (1 GPU: 368^3 Poisson, 272^3 x 3 momentum)
cd "$GRAFT_REPO_ROOT"
timeout 1500 python profiles/run_workloads.py --full > gpurun_out/r02_workloads_full_1gpu.jsonl 2> gpurun_out/r02_workloads_full_1gpu.err
tail -c 1500 gpurun_out/r02_workloads_full_1gpu.err
cut -c1-900 gpurun_out/r02_workloads_full_1gpu.jsonl
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
