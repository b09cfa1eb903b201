#!/bin/bash
# round-2, 8 GPUs: bench (parity leg + timed steps) at N = 8 and N = 4
cd "$GRAFT_REPO_ROOT"
for N in 8 2; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2962$N bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02m${N}_bench.json 2> gpurun_out/r02m${N}_bench.err
tail -c 400 gpurun_out/r02m${N}_bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r02m${N}_bench.json").read().strip().splitlines()[-1])
print("N=${N}", d["ms_per_step"], d["value"], d["setup_phases_ms"], d["iterations"], d["parity"], d["e2e"])
PY
done
