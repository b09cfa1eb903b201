#!/bin/bash
# 2 GPUs, final sources: the NCCL + NVLink peer-memory tests the 1-GPU box skips, then the bench (parity leg + timed steps + e2e)
cd "$GRAFT_REPO_ROOT"
nvidia-smi --query-gpu=index,name --format=csv,noheader; nproc
timeout 900 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu -k "nccl" 2>&1 | tail -8 > gpurun_out/r02n2c_pytest_multirank_nccl.log
cat gpurun_out/r02n2c_pytest_multirank_nccl.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02n2c_bench.json 2> gpurun_out/r02n2c_bench.err
tail -c 400 gpurun_out/r02n2c_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02n2c_bench.json").read().strip().splitlines()[-1])
print("N=2", d["ms_per_step"], d["value"], d["setup_s"], d["solve_s"], d["iterations"], d["parity"]["ok"], d["parity"]["transport"], d["e2e"])
PY
