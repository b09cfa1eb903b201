#!/bin/bash
# round-2 call 1: parity with the flattened interpolation kernel, then A/B bench against the round-1 kernel
set -x
cd "$GRAFT_REPO_ROOT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r02a_pytest_parity.log
cat gpurun_out/r02a_pytest_parity.log
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/r02a_bench_new.json 2> gpurun_out/r02a_bench_new.err
B200_INTERP_V1=1 timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/r02a_bench_v1.json 2> gpurun_out/r02a_bench_v1.err
tail -c 600 gpurun_out/r02a_bench_new.err gpurun_out/r02a_bench_v1.err
python - <<'PY'
import json
for n in ("new","v1"):
    try:
        d=json.loads(open(f"gpurun_out/r02a_bench_{n}.json").read().strip().splitlines()[-1])
        print(n, d["ms_per_step"], d["setup_phases_ms"], d["iterations"], d["level_rows"], {k:(v["ms_total"],v["calls"]) for k,v in d["kernel_classes"].items() if "interp" in k or "spgemm" in k})
    except Exception as e:
        print(n, "failed", e)
PY
