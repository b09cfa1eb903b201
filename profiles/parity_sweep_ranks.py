"""Randomised multi-rank parity sweep: N ranks sharing one GPU (host-staged transport) against the oracle with N virtual ranks.
    python profiles/parity_sweep_ranks.py [cases] [seed] > gpurun_out/parity_sweep_ranks.jsonl
Every case draws a process grid (2 .. 8 ranks), a box per rank, a stencil, the interpolation type, truncation options and the halo
numbering (true neighbours or the reference glue's surrogate), runs the ranks as separate processes (tests/mr_worker.py) and applies
the comparison of tests/test_gpu_multirank.py: levels, CF, S / P / R / A_l patterns bit for bit, values 1e-10, iterations within one.
One JSON line per case and a summary; exit status 1 if any case differs."""
import json
import os
import sys
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 30
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 20261106
    import test_gpu_multirank as T
    rng = np.random.default_rng(seed)
    grids = [(2, 1, 1), (2, 2, 1), (1, 2, 2), (3, 1, 1), (2, 2, 2), (4, 1, 1), (1, 1, 2), (3, 2, 1)]
    bad = 0
    for k in range(cases):
        grid = grids[int(rng.integers(len(grids)))]
        stencil = int(rng.choice([7, 27]))
        n = tuple(int(v) for v in rng.integers(5, 13 if stencil == 7 else 9, size=3))
        theta = float(rng.choice([0.25, 0.5, 0.57]))
        interp = int(rng.choice([6, 6, 0]))
        pmax = int(rng.choice([0, 2, 4, 4]))
        tf = float(rng.choice([0.0, 0.0, 0.2]))
        surrogate = bool(rng.integers(3) == 0)
        relax = int(rng.choice([18, 18, 7, 16, 11, 6, 8]))
        order = int(rng.choice([0, 0, 1])) if relax in (18, 7, 6, 8) else 0
        nloc = n[0] * n[1] * n[2]
        amg = dict(strong_threshold=theta, interp_type=interp, p_max_elmts=pmax, trunc_factor=tf, relax_type=relax, relax_order=order)
        rec = {"case": k, "grid": grid, "n": n, "stencil": stencil, "halo_map": "reference" if surrogate else "true", **amg}
        try:
            cfg = T.settings(n, grid, stencil, **amg)
            extra = dict(interp_type=interp, p_max_elmts=pmax, trunc_factor=tf, relax_type=relax, relax_order=order)
            if surrogate:
                cfg["linear_system"]["halo_map"] = "reference"
                extra["halo_map"] = 1
            res = T.run_ranks(grid[0] * grid[1] * grid[2], cfg)
            T.compare_with_oracle(res, n, grid, stencil, theta, **extra)
            rec.update(ok=True, levels=int(res[0]["num_levels"]), iterations=int(res[0]["iterations"]), rows=nloc * len(res))
        except Exception as e:   # AssertionError of the comparison, or a failed rank
            rec.update(ok=False, why="".join(traceback.format_exception_only(type(e), e))[-600:])
            bad += 1
        print(json.dumps(rec), flush=True)
    print(json.dumps({"summary": True, "cases": cases, "seed": seed, "failed": bad}), flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
