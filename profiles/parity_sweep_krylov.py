"""Randomised sweep of the sibling Krylov methods (FlexGMRES, COGMRES, PCG, BiCGSTAB, GMRES with MGS / CGS2) through the C ABI
against the CPU oracle: random box, stencil, restart length, preconditioner on / off, AMG options.
    python profiles/parity_sweep_krylov.py [cases] [seed] > gpurun_out/parity_sweep_krylov.jsonl
Bars of tests/test_gpu_krylov.py: converged, iterations within one, final relative residual at or below the oracle's, solution to
1e-8 when the counts agree.  One JSON line per case and a summary; exit status 1 if any case differs."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
TOL = 1e-8


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 80
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    import b200hypre as bh
    from oracle_lib import Oracle
    bh.comm_init(0, 1)
    rng = np.random.default_rng(seed)
    bad = 0
    for k in range(cases):
        method = str(rng.choice(["fgmres", "cogmres", "cg", "bicg", "gmres"]))
        precond = int(rng.integers(4) != 0)
        stencil = int(rng.choice([7, 27]))
        n = tuple(int(v) for v in rng.integers(5, 24 if precond else 13, size=3))
        kspace = int(rng.choice([5, 10, 20, 50]))
        interp = int(rng.choice([6, 6, 0]))
        theta = float(rng.choice([0.25, 0.57]))
        relax = int(rng.choice([18, 7, 11]))
        amg = dict(strong_threshold=theta, interp_type=interp, relax_type=relax, relax_order=0, coarsen_type=8) if precond else {}
        extra = {}
        if method == "cogmres":
            extra["cgs"] = int(rng.choice([0, 1, 2]))
        if method == "gmres":
            extra["orthogonalization"] = "mgs"
        rec = {"case": k, "method": method, "precond": precond, "n": n, "stencil": stencil, "kspace": kspace, **amg, **extra}
        s = bh.HypreSystem({"linear_system": {"type": f"build_{stencil}pt_stencil", "nx": n[0], "ny": n[1], "nz": n[2]},
                            "solver_settings": dict({"method": method, "preconditioner": "boomeramg" if precond else "none", "tolerance": TOL,
                                                     "kspace": kspace, "max_iterations": 600}, **extra),
                            "boomeramg_settings": amg})
        s.setup_precon_and_solver()
        s.load()
        s.solve()
        oopts = dict(tolerance=TOL, kspace=kspace, max_iterations=600, preconditioner=precond)
        if precond:
            oopts.update(strong_threshold=theta, interp_type=interp, relax_type=relax, relax_order=0)
        if "cgs" in extra:
            oopts["cgs"] = extra["cgs"]
        o = Oracle(**oopts)
        o.threads(len(os.sched_getaffinity(0)))
        o.gen_laplace(n, stencil=stencil)
        if precond:
            o.setup()
        xo = o.solve(method=method)
        x = s.solution()
        why = []
        if s.solver.converged != 1:
            why.append("not converged")
        if abs(s.solver.iterations - o.iterations) > 1:
            why.append("iterations")
        if s.solver.relres > max(o.relres * (1 + 1e-6), TOL):
            why.append("relres")
        if s.solver.iterations == o.iterations and np.abs(x - xo).max() / np.abs(xo).max() > 1e-8:
            why.append("solution")
        if np.abs(x - 1.0).max() > 1e-5:
            why.append("known answer")
        rec.update(iters=[int(s.solver.iterations), int(o.iterations)], relres=[float(s.solver.relres), float(o.relres)], ok=not why, why=why)
        bad += 0 if rec["ok"] else 1
        print(json.dumps(rec), flush=True)
        s.destroy_system()
    print(json.dumps({"summary": True, "cases": cases, "seed": seed, "failed": bad}), flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
