#!/bin/bash
# round-2: configs[4] at its stated size (322^3 x 3 = 100.2 M rows, 2.10 G non-zeros) on 2 GPUs
cd "$GRAFT_REPO_ROOT"
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29641 profiles/run_workloads.py --full > gpurun_out/r02_workloads_full_2gpu.jsonl 2> gpurun_out/r02_workloads_full_2gpu.err
tail -c 1500 gpurun_out/r02_workloads_full_2gpu.err
cut -c1-900 gpurun_out/r02_workloads_full_2gpu.jsonl
