"""Randomised parity sweep, GPU path (through the C ABI) against the CPU oracle -- more cases than the test suite can afford:
    python profiles/parity_sweep.py [cases] [seed] > gpurun_out/parity_sweep.jsonl
Every case draws a box (anisotropic sizes allowed), a stencil, a strength threshold, the interpolation type (extended+i or classical),
P_max_elmts / trunc_factor and a smoother, builds the hierarchy on the device and in the oracle, and compares: number of levels, CF
splitting, row pointers and column order of every A_l and P_l bit for bit, values to 1e-10, GMRES iterations within one, solution
against the generator's known answer.  One JSON line per case, a summary line at the end; exit status 1 if any case differs."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def one_case(bh, Oracle, rng, k):
    stencil = int(rng.choice([7, 27]))
    hi = 36 if stencil == 7 else 22
    n = tuple(int(v) for v in rng.integers(5, hi, size=3))
    theta = float(rng.choice([0.25, 0.5, 0.57, 0.7]))
    interp = int(rng.choice([6, 6, 0]))
    pmax = int(rng.choice([0, 2, 4, 4, 6]))
    tf = float(rng.choice([0.0, 0.0, 0.1, 0.3]))
    relax = int(rng.choice([18, 18, 7, 16, 11, 12, 6, 8, 3]))
    order = int(rng.choice([0, 0, 1])) if relax in (18, 7, 6, 8, 3) else 0
    amg = dict(strong_threshold=theta, interp_type=interp, p_max_elmts=pmax, trunc_factor=tf, relax_type=relax, relax_order=order, coarsen_type=8)
    rec = {"case": k, "n": n, "stencil": stencil, **amg}
    s = bh.HypreSystem({"linear_system": {"type": f"build_{stencil}pt_stencil", "nx": n[0], "ny": n[1], "nz": n[2]},
                        "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50,
                                            "max_iterations": 300, "orthogonalization": "mgs"},
                        "boomeramg_settings": amg})
    s.setup_precon_and_solver()
    s.load()
    s.solve()
    o = Oracle(tolerance=1e-8, kspace=50, max_iterations=300, strong_threshold=theta, interp_type=interp, p_max_elmts=pmax,
               trunc_factor=tf, relax_type=relax, relax_order=order)
    o.threads(len(os.sched_getaffinity(0)))
    o.gen_laplace(n, stencil=stencil)
    o.setup()
    o.solve()
    g = s.precond
    why = []
    if g.num_levels != o.num_levels:
        why.append(f"levels {g.num_levels} vs {o.num_levels}")
    else:
        worst = 0.0
        for l in range(o.num_levels):
            for which in (("A", "P") if l + 1 < o.num_levels else ("A",)):
                gp, gc, gv = g.level_matrix(l, which)
                rp, rc, rv, _ = o.matrix(l, which)
                if not (np.array_equal(gp, rp) and np.array_equal(gc, rc.astype(gc.dtype))):
                    why.append(f"{which}[{l}] pattern")
                elif len(rv):
                    worst = max(worst, float(np.abs(gv - rv).max() / np.abs(rv).max()))
            if l + 1 < o.num_levels and not np.array_equal(g.level_cf(l), o.cf(l)):
                why.append(f"CF[{l}]")
        rec["val_relerr"] = worst
        if worst > 1e-10:
            why.append(f"values {worst:.2e}")
    rec.update(levels=int(g.num_levels), iters=[int(s.solver.iterations), int(o.iterations)], converged=int(s.solver.converged),
               err_vs_1=float(np.abs(s.solution() - 1.0).max()))
    if abs(s.solver.iterations - o.iterations) > 1:
        why.append("iterations")
    if s.solver.converged != 1 or rec["err_vs_1"] > 1e-5:
        why.append("solve")
    rec["ok"] = not why
    rec["why"] = why
    s.destroy_system()
    return rec


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 20261105
    import b200hypre as bh
    from oracle_lib import Oracle
    bh.comm_init(0, 1)
    rng = np.random.default_rng(seed)
    bad = 0
    for k in range(cases):
        rec = one_case(bh, Oracle, rng, k)
        bad += 0 if rec["ok"] else 1
        print(json.dumps(rec), flush=True)
    print(json.dumps({"summary": True, "cases": cases, "seed": seed, "failed": bad}), flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
