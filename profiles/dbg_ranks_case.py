"""re-run one multi-rank configuration and print every rank's output on failure:
    python profiles/dbg_ranks_case.py '<grid>' '<n>' stencil theta interp pmax tf [reference]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_gpu_multirank as T
grid, n = tuple(json.loads(sys.argv[1])), tuple(json.loads(sys.argv[2]))
stencil, theta, interp, pmax, tf = int(sys.argv[3]), float(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6]), float(sys.argv[7])
cfg = T.settings(n, grid, stencil, strong_threshold=theta, interp_type=interp, p_max_elmts=pmax, trunc_factor=tf)
extra = dict(interp_type=interp, p_max_elmts=pmax, trunc_factor=tf)
if len(sys.argv) > 8:
    cfg["linear_system"]["halo_map"] = "reference"; extra["halo_map"] = 1
try:
    res = T.run_ranks(grid[0] * grid[1] * grid[2], cfg)
    T.compare_with_oracle(res, n, grid, stencil, theta, **extra)
    print("OK", int(res[0]["num_levels"]), int(res[0]["iterations"]))
except AssertionError as e:
    print(str(e)[-12000:])
