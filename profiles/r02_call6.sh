#!/bin/bash
# round-2 call 6: whole GPU suite on the C-compressed interpolation kernel + fused PMIS sweep, then phase timing, bench, ncu
cd "$GRAFT_REPO_ROOT"
timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r02f_pytest_gpu.log
cat gpurun_out/r02f_pytest_gpu.log
B200_INTERP_TIMING=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --parity-n 0 > gpurun_out/r02f_timing.json 2> gpurun_out/r02f_timing.err
grep "interp timing" gpurun_out/r02f_timing.err | tail -9
timeout 600 python bench.py --steps 5 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/r02f_bench.json 2> gpurun_out/r02f_bench.err
tail -c 400 gpurun_out/r02f_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02f_bench.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["setup_phases_ms"], d["iterations"], d["parity"]["ok"], "pool", d["pool_high_water_gib"])
print({k:(v["ms_total"],v["calls"]) for k,v in d["kernel_classes"].items() if "interp" in k or "spgemm" in k})
PY
python bench.py --size 256 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --parity-n 0 > gpurun_out/r02f_plain.log 2>&1 &&
bash profiles/prof_kernel.sh r02f_interpL0 'k_interp_flat' 0 1 --parity-n 0 &&
bash profiles/prof_kernel.sh r02f_interpL1 'k_interp_flat<\(int\)32' 0 1 --parity-n 0 &&
bash profiles/prof_kernel.sh r02f_trunc 'k_interp_trunc' 0 1 --parity-n 0
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r02f_traffic.csv -k regex:'k_interp_flat|k_interp_trunc|k_interp_ovf|k_spgemm' python bench.py --size 256 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --parity-n 0 > gpurun_out/r02f_traffic.log 2>&1
tail -2 gpurun_out/r02f_traffic.log | cut -c1-200
