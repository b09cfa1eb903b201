#!/bin/bash
# round-2: ncu --set full of the first SpGEMM launches of one step (A*P and R*Q of level 0: symbolic and numeric bins)
cd "$GRAFT_REPO_ROOT"
python bench.py --size 256 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --parity-n 0 > gpurun_out/r02g_plain.log 2>&1 &&
bash profiles/prof_kernel.sh r02g_spgemm 'k_spgemm_smem|k_spgemm_count_try' 0 10 --parity-n 0
