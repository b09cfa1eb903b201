#!/bin/bash
# round-2, 2 GPUs: the NCCL + NVLink peer-memory tests the 1-GPU box skips, then the bench (parity leg + timed steps) and the
# reference arm at N = 2
cd "$GRAFT_REPO_ROOT"
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 1500 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu 2>&1 | tail -8 > gpurun_out/r02n2_pytest_multirank.log
cat gpurun_out/r02n2_pytest_multirank.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02n2_bench.json 2> gpurun_out/r02n2_bench.err
tail -c 600 gpurun_out/r02n2_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02n2_bench.json").read().strip().splitlines()[-1])
print("N=2", d["ms_per_step"], d["value"], d["setup_phases_ms"], d["iterations"], d["parity"], d["e2e"])
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --impl reference --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02n2_reference.json 2> gpurun_out/r02n2_reference.err
tail -c 300 gpurun_out/r02n2_reference.err; cut -c1-700 gpurun_out/r02n2_reference.json
