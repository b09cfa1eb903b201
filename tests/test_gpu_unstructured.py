"""GPU parity on UNSTRUCTURED matrices (every other parity test uses grid stencils or the two grid-based workloads): weighted
k-nearest-neighbour graphs of random points (tests/workloads.py: knn_graph_matrix) with hub rows of hundreds of entries and a fraction of
positive off-diagonals, through the IJ interface as CSR; one rank, and block rows over several ranks sharing the GPU -- the points are
in random order, so nearly every connection leaves its rank.  Bars as everywhere: hierarchy bit for bit, values 1e-10, iterations +-1.
profiles/parity_sweep_unstructured.py runs many more of these."""
import os
import tempfile

import numpy as np
import pytest

from oracle_lib import Oracle
from workloads import knn_graph_matrix
import test_gpu_multirank as T

pytestmark = pytest.mark.gpu
VAL_RTOL = 1e-10


@pytest.mark.parametrize("n,dim,k,hubs,pos,interp,pmax,tf,relax", [
    (9000, 3, 8, 2, 0.1, 6, 4, 0.0, 18),      # hub rows of ~700 entries: beyond the 32-lane tier's C-hat capacity on level 0
    (7000, 2, 16, 0, 0.0, 6, 0, 0.2, 7),      # long regular rows, trunc_factor only
    (8000, 3, 32, 1, 0.1, 0, 4, 0.0, 11),     # classical interpolation, two-stage Gauss-Seidel
    (5000, 2, 4, 3, 0.0, 0, 0, 0.0, 18),
])
def test_unstructured_hierarchy_bit_exact(b200, n, dim, k, hubs, pos, interp, pmax, tf, relax):
    rng = np.random.default_rng(n + k)
    A = knn_graph_matrix(rng, n, dim, k, hubs, 700, pos, 0.01)
    b = A @ np.ones(n)
    amg = dict(strong_threshold=0.25, interp_type=interp, p_max_elmts=pmax, trunc_factor=tf, relax_type=relax, relax_order=0, coarsen_type=8)
    s = b200.HypreSystem({"linear_system": {"type": "csr", "ptr": A.indptr, "col": A.indices, "val": A.data, "rhs": b},
                          "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50,
                                              "max_iterations": 400, "orthogonalization": "mgs"},
                          "boomeramg_settings": amg})
    s.setup_precon_and_solver()
    s.load()
    s.solve()
    o = Oracle(tolerance=1e-8, kspace=50, max_iterations=400, strong_threshold=0.25, interp_type=interp, p_max_elmts=pmax, trunc_factor=tf,
               relax_type=relax, relax_order=0)
    o.threads(8)
    o.set_csr(A.indptr, A.indices, A.data, b)
    o.setup()
    o.solve()
    g = s.precond
    assert g.num_levels == o.num_levels
    for l in range(o.num_levels):
        for which in (("A", "P") if l + 1 < o.num_levels else ("A",)):
            gp, gc, gv = g.level_matrix(l, which)
            rp, rc, rv, _ = o.matrix(l, which)
            assert np.array_equal(gp, rp) and np.array_equal(gc, rc.astype(gc.dtype)), f"{which}[{l}] pattern"
            if len(rv):
                assert np.abs(gv - rv).max() / np.abs(rv).max() <= VAL_RTOL, f"{which}[{l}] values"
        if l + 1 < o.num_levels:
            assert np.array_equal(g.level_cf(l), o.cf(l)), f"CF[{l}]"
    assert abs(s.solver.iterations - o.iterations) <= 1
    assert s.solver.converged == 1 and np.abs(s.solution() - 1.0).max() < 1e-5
    s.destroy_system()


@pytest.mark.parametrize("nranks,n,k,interp", [(3, 2500, 8, 6), (4, 3000, 4, 0), (8, 2000, 16, 6)])
def test_unstructured_block_rows_over_ranks(b200, nranks, n, k, interp):
    rng = np.random.default_rng(100 * nranks + k)
    A = knn_graph_matrix(rng, n, 3, k, 2, 150, 0.1, 0.01)
    b = A @ np.ones(n)
    amg = dict(strong_threshold=0.25, interp_type=interp, p_max_elmts=4, trunc_factor=0.0, relax_type=18, relax_order=0)
    with tempfile.TemporaryDirectory() as tmp:
        f = os.path.join(tmp, "sys.npz")
        np.savez(f, ptr=A.indptr, col=A.indices, val=A.data, rhs=b)
        cfg = {"linear_system": {"type": "csr_npz", "file": f},
               "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 400},
               "boomeramg_settings": amg}
        res = T.run_ranks(nranks, cfg)
    o = Oracle(tolerance=1e-8, kspace=50, max_iterations=400, **amg)
    o.threads(8)
    o.set_csr(A.indptr, A.indices, A.data, b, nparts=nranks)
    o.setup()
    o.solve()
    nl = int(res[0]["num_levels"])
    assert nl == o.num_levels
    for l in range(nl):
        T.check_csr(T.stack_csr(res, f"A{l}"), o.matrix(l, "A"), f"A[{l}]")
        if l + 1 < nl:
            assert np.array_equal(np.concatenate([d[f"cf{l}"] for d in res]), o.cf(l)), f"CF[{l}]"
            T.check_csr(T.stack_csr(res, f"P{l}"), o.matrix(l, "P"), f"P[{l}]")
            T.check_csr(T.stack_csr(res, f"R{l}"), o.matrix(l, "R"), f"R[{l}]")
    assert abs(int(res[0]["iterations"]) - o.iterations) <= 1
    assert np.abs(np.concatenate([d["x"] for d in res]) - 1.0).max() < 1e-5
