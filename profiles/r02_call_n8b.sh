#!/bin/bash
# 8 GPUs, final sources: bench (parity leg + timed steps + e2e)
cd "$GRAFT_REPO_ROOT"
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8; nproc
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r02n8c_bench.json 2> gpurun_out/r02n8c_bench.err
tail -c 400 gpurun_out/r02n8c_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02n8c_bench.json").read().strip().splitlines()[-1])
print("N=8", d["ms_per_step"], d["value"], d["setup_s"], d["solve_s"], d["iterations"], d["parity"]["ok"], d["parity"]["transport"], d["e2e"])
PY
