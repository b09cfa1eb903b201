#!/bin/bash
# final sources of round 2: the allocation-free test alone in a fresh process, the whole GPU suite (that file is collected last), the traffic
# capture of this csrc/ hash, the stamp, and the default bench line that carries it
cd "$GRAFT_REPO_ROOT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader; nproc
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -8 > gpurun_out/r02w_pytest_noalloc_alone.log
cat gpurun_out/r02w_pytest_noalloc_alone.log
B200_SHM_TIMEOUT_S=60 timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -8 > gpurun_out/r02w_pytest_gpu.log
cat gpurun_out/r02w_pytest_gpu.log
timeout 600 python bench.py --size 256 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --parity-n 0 > gpurun_out/r02w_plain.log 2>&1 || { tail -5 gpurun_out/r02w_plain.log; exit 1; }
timeout 1200 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r02w_traffic.csv -k regex:'k_interp_flat|k_interp_trunc|k_interp_ovf|k_spgemm' python bench.py --size 256 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --parity-n 0 > gpurun_out/r02w_traffic.log 2>&1
python profiles/make_roofline_traffic.py gpurun_out/r02w_traffic.csv && cp profiles/roofline_traffic.json gpurun_out/r02w_roofline_traffic.json
timeout 900 python bench.py > gpurun_out/r02w_bench_256.json 2> gpurun_out/r02w_bench_256.err
tail -c 300 gpurun_out/r02w_bench_256.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02w_bench_256.json").read().strip().splitlines()[-1])
print("b200", d["ms_per_step"], d["value"], "e2e", d["e2e"]["ms_per_step"], d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], "parity", d["parity"]["ok"], "roofline", d["roofline"]["kernel"], d["roofline"]["frac"], d["roofline"]["traffic"])
PY
