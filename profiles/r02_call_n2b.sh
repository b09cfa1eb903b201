#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29615 bench.py --gpus 2 --steps 4 --warmup 2 --no-e2e --parity-n 0 > gpurun_out/r02n2b_bench.json 2> gpurun_out/r02n2b_bench.err
tail -c 300 gpurun_out/r02n2b_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02n2b_bench.json").read().strip().splitlines()[-1])
print("N=2", d["ms_per_step"], "setup", d["setup_s"], "solve", d["solve_s"])
for k,v in sorted(d["setup_phases_ms"].items()): print(f"  {k:22s} {v:8.2f}")
PY
timeout 600 python bench.py --steps 4 --warmup 2 --no-e2e --no-cpu-baseline --parity-n 0 > gpurun_out/r02n1b_bench.json 2> gpurun_out/r02n1b_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02n1b_bench.json").read().strip().splitlines()[-1])
print("N=1", d["ms_per_step"], "setup", d["setup_s"], "solve", d["solve_s"])
for k,v in sorted(d["setup_phases_ms"].items()): print(f"  {k:22s} {v:8.2f}")
PY
