"""Randomised parity sweep on UNSTRUCTURED matrices (the suite and the other sweeps use grid stencils): weighted k-nearest-neighbour
graphs of random points in 2-D / 3-D, symmetrised, optionally with a few hub rows of hundreds of entries (the capacity tiers and the
overflow paths of the interpolation and SpGEMM kernels) and with a fraction of positive off-diagonals (the sign rules of the
interpolation formulas), made diagonally dominant; rhs = A 1.  The matrices enter through the IJ interface as CSR.
    python profiles/parity_sweep_unstructured.py [cases] [seed] > gpurun_out/parity_sweep_unstructured.jsonl
Comparison as in profiles/parity_sweep.py: levels, CF, patterns of A_l / P_l bit for bit, values 1e-10, iterations within one."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def random_matrix(rng, n, dim, k, hubs, hub_deg, positive_frac, shift):
    from workloads import knn_graph_matrix   # tests/workloads.py: the generator the test suite uses as well
    return knn_graph_matrix(rng, n, dim, k, hubs, hub_deg, positive_frac, shift)


def main_ranks(cases, seed):
    """--ranks: the same matrices, block rows over 2 .. 8 ranks sharing the GPU (tests/mr_worker.py), against the N-virtual-rank
    oracle.  The points are in random order, so nearly every connection of a row leaves its rank: many halo points, every rank a
    neighbour of every other, distance-two points everywhere."""
    import tempfile
    import traceback
    import test_gpu_multirank as T
    from oracle_lib import Oracle
    rng = np.random.default_rng(seed)
    bad = 0
    for c in range(cases):
        nranks = int(rng.choice([2, 3, 4, 8]))
        n = int(rng.integers(800, 6000))
        dim = int(rng.choice([2, 3]))
        k = int(rng.choice([4, 8, 16]))
        hubs = int(rng.choice([0, 0, 2]))
        pos = float(rng.choice([0.0, 0.0, 0.1]))
        shift = float(rng.choice([0.0, 0.01, 0.2]))
        theta = float(rng.choice([0.25, 0.5]))
        interp = int(rng.choice([6, 6, 0]))
        pmax = int(rng.choice([0, 4, 4]))
        tf = float(rng.choice([0.0, 0.0, 0.2]))
        relax = int(rng.choice([18, 18, 7, 11, 16]))
        A = random_matrix(rng, n, dim, k, hubs, 150, pos, shift)
        b = A @ np.ones(n)
        amg = dict(strong_threshold=theta, interp_type=interp, p_max_elmts=pmax, trunc_factor=tf, relax_type=relax, relax_order=0)
        rec = {"case": c, "ranks": nranks, "n": n, "dim": dim, "k": k, "hubs": hubs, "positive_frac": pos, "shift": shift, "nnz": int(A.nnz), **amg}
        try:
            with tempfile.TemporaryDirectory() as tmp:
                f = os.path.join(tmp, "sys.npz")
                np.savez(f, ptr=A.indptr, col=A.indices, val=A.data, rhs=b)
                cfg = {"linear_system": {"type": "csr_npz", "file": f},
                       "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 400},
                       "boomeramg_settings": amg}
                res = T.run_ranks(nranks, cfg)
            o = Oracle(tolerance=1e-8, kspace=50, max_iterations=400, **amg)
            o.threads(len(os.sched_getaffinity(0)))
            o.set_csr(A.indptr, A.indices, A.data, b, nparts=nranks)
            o.setup()
            xo = o.solve()
            nl = int(res[0]["num_levels"])
            assert nl == o.num_levels, (nl, o.num_levels)
            for l in range(nl):
                T.check_csr(T.stack_csr(res, f"A{l}"), o.matrix(l, "A"), f"A[{l}]")
                if l + 1 < nl:
                    cf = np.concatenate([d[f"cf{l}"] for d in res])
                    assert np.array_equal(cf, o.cf(l)), f"CF[{l}] differs"
                    T.check_csr(T.stack_csr(res, f"P{l}"), o.matrix(l, "P"), f"P[{l}]")
                    T.check_csr(T.stack_csr(res, f"R{l}"), o.matrix(l, "R"), f"R[{l}]")
            its = int(res[0]["iterations"])
            assert abs(its - o.iterations) <= 1, (its, o.iterations)
            x = np.concatenate([d["x"] for d in res])
            assert np.abs(x - 1.0).max() < 1e-5
            rec.update(ok=True, levels=nl, iters=[its, int(o.iterations)])
        except Exception as e:
            rec.update(ok=False, why="".join(traceback.format_exception_only(type(e), e))[-1500:])
            bad += 1
        print(json.dumps(rec), flush=True)
    print(json.dumps({"summary": True, "ranks": True, "cases": cases, "seed": seed, "failed": bad}), flush=True)
    sys.exit(1 if bad else 0)


def main():
    if "--ranks" in sys.argv:
        args = [a for a in sys.argv[1:] if a != "--ranks"]
        return main_ranks(int(args[0]) if args else 30, int(args[1]) if len(args) > 1 else 4)
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    import b200hypre as bh
    from oracle_lib import Oracle
    bh.comm_init(0, 1)
    rng = np.random.default_rng(seed)
    bad = 0
    for c in range(cases):
        n = int(rng.integers(1500, 25000))
        dim = int(rng.choice([2, 3]))
        k = int(rng.choice([4, 8, 16, 32]))
        hubs = int(rng.choice([0, 0, 1, 3]))
        hub_deg = int(rng.choice([120, 300, 700]))
        pos = float(rng.choice([0.0, 0.0, 0.1]))
        shift = float(rng.choice([0.0, 0.01, 0.2]))
        theta = float(rng.choice([0.25, 0.5]))
        interp = int(rng.choice([6, 6, 0]))
        pmax = int(rng.choice([0, 4, 4]))
        tf = float(rng.choice([0.0, 0.0, 0.2]))
        relax = int(rng.choice([18, 18, 7, 11]))
        A = random_matrix(rng, n, dim, k, hubs, hub_deg, pos, shift)
        b = A @ np.ones(n)
        amg = dict(strong_threshold=theta, interp_type=interp, p_max_elmts=pmax, trunc_factor=tf, relax_type=relax, relax_order=0, coarsen_type=8)
        rec = {"case": c, "n": n, "dim": dim, "k": k, "hubs": hubs, "hub_deg": hub_deg if hubs else 0, "positive_frac": pos, "shift": shift,
               "nnz": int(A.nnz), "max_row": int(np.diff(A.indptr).max()), **amg}
        s = bh.HypreSystem({"linear_system": {"type": "csr", "ptr": A.indptr, "col": A.indices, "val": A.data, "rhs": b},
                            "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50,
                                                "max_iterations": 400, "orthogonalization": "mgs"},
                            "boomeramg_settings": amg})
        s.setup_precon_and_solver()
        s.load()
        s.solve()
        o = Oracle(tolerance=1e-8, kspace=50, max_iterations=400, strong_threshold=theta, interp_type=interp, p_max_elmts=pmax,
                   trunc_factor=tf, relax_type=relax, relax_order=0)
        o.threads(len(os.sched_getaffinity(0)))
        o.set_csr(A.indptr, A.indices, A.data, b)
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
        if s.solver.converged != 1 and o.iterations < 400:
            why.append("solve")
        rec["ok"] = not why
        rec["why"] = why
        bad += 0 if rec["ok"] else 1
        print(json.dumps(rec), flush=True)
        s.destroy_system()
    print(json.dumps({"summary": True, "cases": cases, "seed": seed, "failed": bad}), flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
