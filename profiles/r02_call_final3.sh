#!/bin/bash
# round-2 final: GPU suite, default bench line (cpu_baseline + e2e + parity leg), reference arm, launch list and DRAM traffic
cd "$GRAFT_REPO_ROOT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader; nproc
timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 > gpurun_out/r02x_pytest_gpu.log
cat gpurun_out/r02x_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/r02x_bench_256.json 2> gpurun_out/r02x_bench_256.err
tail -c 300 gpurun_out/r02x_bench_256.err
timeout 900 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r02x_bench_reference.json 2> gpurun_out/r02x_bench_reference.err
tail -c 300 gpurun_out/r02x_bench_reference.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02x_bench_256.json").read().strip().splitlines()[-1])
print("b200", d["ms_per_step"], d["value"], "e2e", d["e2e"]["ms_per_step"], d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], "parity", d["parity"]["ok"], "roofline", d["roofline"]["kernel"], d["roofline"]["frac"])
r=json.loads(open("gpurun_out/r02x_bench_reference.json").read().strip().splitlines()[-1])
print("reference", r["ms_per_step"], r["value"], r["cpu_baseline"]["cores"], r["sample"][:200])
PY
bash profiles/prof_step.sh r02x > gpurun_out/r02x_prof_step.log 2>&1
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r02x_traffic.csv -k regex:'k_interp_flat|k_interp_trunc|k_interp_ovf|k_spgemm' python bench.py --size 256 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --parity-n 0 > gpurun_out/r02x_traffic.log 2>&1
ls -la gpurun_out/launches_r02x.csv gpurun_out/r02x_traffic.csv
