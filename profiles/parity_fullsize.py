"""Hierarchy parity at the headline size (run under gpurun): BoomerAMG setup of the 7-pt Laplacian with n^3 rows on the GPU
(through the C ABI) and on the CPU oracle, compared level by level -- sizes, CF splitting, patterns of A / P bit-for-bit,
values by maximum relative difference -- plus iteration counts and solutions of the GMRES solve.
    python profiles/parity_fullsize.py 256 > gpurun_out/parity_256.jsonl            # 7-pt, theta 0.25
    python profiles/parity_fullsize.py 160 27 0.57 > gpurun_out/parity_27pt_160.jsonl
    python profiles/parity_fullsize.py 256 7 0.25 0 > gpurun_out/parity_256_classical.jsonl   # interp_type 0
The oracle is test infrastructure; this script is a checker like the tests, kept out of the -m gpu suite because of its size."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    stencil = int(sys.argv[2]) if len(sys.argv) > 2 else 7
    theta = float(sys.argv[3]) if len(sys.argv) > 3 else 0.25
    interp = int(sys.argv[4]) if len(sys.argv) > 4 else 6      # 6 = extended+i, 0 = classical modified
    import b200hypre as bh
    from oracle_lib import Oracle
    bh.comm_init(0, 1)
    s = bh.HypreSystem({"linear_system": {"type": f"build_{stencil}pt_stencil", "nx": n, "ny": n, "nz": n},
                        "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50,
                                            "max_iterations": 200, "orthogonalization": "mgs"},
                        "boomeramg_settings": {"strong_threshold": theta, "interp_type": interp}})
    s.setup_precon_and_solver()
    s.load()
    s.solve()
    t0 = time.time()
    o = Oracle(tolerance=1e-8, kspace=50, max_iterations=200, strong_threshold=theta, interp_type=interp)
    cores = o.threads(0)
    o.gen_laplace(n, stencil=stencil)
    o.setup()
    xo = o.solve()
    t_oracle = time.time() - t0
    amg = s.precond
    ok = amg.num_levels == o.num_levels
    print(json.dumps({"n": n, "stencil": stencil, "strong_threshold": theta, "rows": n ** 3, "levels_gpu": amg.num_levels, "levels_oracle": o.num_levels, "oracle_seconds": round(t_oracle, 1),
                      "oracle_threads": cores}), flush=True)
    for l in range(min(amg.num_levels, o.num_levels)):
        rec = {"level": l}
        for which in ("A", "P") if l + 1 < o.num_levels else ("A",):
            gp, gc, gv = amg.level_matrix(l, which)
            rp, rc, rv, shape = o.matrix(l, which)
            same_ptr = bool(np.array_equal(gp, rp))
            same_col = bool(same_ptr and np.array_equal(gc, rc.astype(gc.dtype)))
            err = float(np.abs(gv - rv).max() / np.abs(rv).max()) if same_col and len(rv) else None
            rec[which] = {"rows": int(shape[0]), "nnz": int(len(rc)), "pattern_identical": same_col, "max_rel_value_diff": err}
            ok = ok and same_col and (err is None or err <= 1e-10)
            del gp, gc, gv, rp, rc, rv
        if l + 1 < o.num_levels:
            rec["cf_identical"] = bool(np.array_equal(amg.level_cf(l), o.cf(l)))
            ok = ok and rec["cf_identical"]
        print(json.dumps(rec), flush=True)
    x = s.solution()
    res = {"iterations_gpu": s.solver.iterations, "iterations_oracle": o.iterations, "relres_gpu": s.solver.relres, "relres_oracle": o.relres,
           "max_rel_solution_diff": float(np.abs(x - xo).max() / np.abs(xo).max()), "max_err_vs_exact": float(np.abs(x - 1.0).max())}
    ok = ok and abs(res["iterations_gpu"] - res["iterations_oracle"]) <= 1
    res["parity"] = bool(ok)
    print(json.dumps(res), flush=True)
    s.destroy_system()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
