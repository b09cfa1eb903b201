"""GPU parity tests of the hybrid Gauss-Seidel smoothers (relax_type 3, 4, 6, 8, 13, 14): the
mini-app's default smoother is 8 with relax_order 1 (/root/reference/src/HypreSystem.cpp:138,156).  The wavefront
kernels must reproduce the SEQUENTIAL sweep of the CPU algorithm: one V-cycle agrees with the oracle to 1e-10 and the
preconditioned GMRES takes the same number of iterations (+-1)."""
import numpy as np
import pytest

from oracle_lib import Oracle
from test_gpu_multirank import run_ranks, settings
from workloads import momentum3, poisson27_hex

pytestmark = pytest.mark.gpu

VAL_RTOL = 1e-10


def gen_system(b200, n, stencil, **amg):
    cfg = {
        "linear_system": {"type": "build_27pt_stencil" if stencil == 27 else "build_7pt_stencil", "nx": n, "ny": n, "nz": n},
        "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 200,
                            "orthogonalization": "mgs"},
        "boomeramg_settings": amg,
    }
    s = b200.HypreSystem(cfg)
    s.setup_precon_and_solver()
    s.load()
    return s


def csr_system(b200, A, b, **amg):
    cfg = {
        "linear_system": {"type": "csr", "ptr": A.indptr, "col": A.indices, "val": A.data, "rhs": b},
        "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 400,
                            "orthogonalization": "mgs"},
        "boomeramg_settings": amg,
    }
    s = b200.HypreSystem(cfg)
    s.setup_precon_and_solver()
    s.load()
    return s


def vcycle_gpu(b200, s, r):
    import torch
    rd = torch.tensor(r, device="cuda")
    zd = torch.zeros_like(rd)
    torch.cuda.synchronize()
    s.precond.vcycle_ptr(rd.data_ptr(), zd.data_ptr())
    b200.synchronize()
    return zd.cpu().numpy()


@pytest.mark.parametrize("relax", [3, 4, 6, 8, 13, 14])
@pytest.mark.parametrize("order", [0, 1])
def test_vcycle_and_solve_match_oracle(b200, relax, order):
    n = 20
    s = gen_system(b200, n, 7, relax_type=relax, relax_order=order)
    s.solve()
    o = Oracle(relax_type=relax, relax_order=order)
    o.threads(8)
    o.gen_laplace(n, stencil=7)
    o.setup()
    xo = o.solve()
    r = np.random.default_rng(11).standard_normal(n ** 3)
    z, zo = vcycle_gpu(b200, s, r), o.vcycle(r)
    assert np.abs(z - zo).max() / np.abs(zo).max() <= VAL_RTOL
    assert abs(s.solver.iterations - o.iterations) <= 1, (s.solver.iterations, o.iterations)
    assert s.solver.converged == 1
    assert s.solver.relres <= max(o.relres * (1 + 1e-6), 1e-8)
    x = s.solution()
    assert np.abs(x - 1.0).max() < 1e-6
    if s.solver.iterations == o.iterations:
        assert np.abs(x - xo).max() / np.abs(xo).max() <= 1e-9
    s.destroy_system()


def test_default_smoother_27pt(b200):
    """the mini-app defaults: relax_type 8, relax_order 1, strong_threshold 0.57 on the 27-point generator"""
    n = 14
    s = gen_system(b200, n, 27, relax_type=8, relax_order=1, strong_threshold=0.57)
    s.solve()
    o = Oracle(relax_type=8, relax_order=1, strong_threshold=0.57)
    o.threads(8)
    o.gen_laplace(n, stencil=27)
    o.setup()
    o.solve()
    r = np.random.default_rng(5).standard_normal(n ** 3)
    z, zo = vcycle_gpu(b200, s, r), o.vcycle(r)
    assert np.abs(z - zo).max() / np.abs(zo).max() <= VAL_RTOL
    assert abs(s.solver.iterations - o.iterations) <= 1
    s.destroy_system()


@pytest.mark.parametrize("relax,order", [(6, 0), (8, 1), (3, 1)])
def test_nonsymmetric_pattern_independent_workloads(b200, relax, order):
    """perturbed-hex Poisson (symmetric) and the 3-DOF momentum system (nonsymmetric values, smoother-only SGS
    preconditioner = BASELINE configs[4] 'SGS')"""
    A, b = poisson27_hex(12)
    s = csr_system(b200, A, b, relax_type=relax, relax_order=order)
    s.solve()
    o = Oracle(relax_type=relax, relax_order=order, tolerance=1e-8, kspace=50, max_iterations=400)
    o.threads(8)
    o.set_csr(A.indptr, A.indices, A.data, b)
    o.setup()
    o.solve()
    r = np.random.default_rng(3).standard_normal(A.shape[0])
    z, zo = vcycle_gpu(b200, s, r), o.vcycle(r)
    assert np.abs(z - zo).max() / np.abs(zo).max() <= VAL_RTOL
    assert abs(s.solver.iterations - o.iterations) <= 1
    s.destroy_system()

    A, b = momentum3(8)
    s = csr_system(b200, A, b, relax_type=relax, relax_order=0, max_levels=1)
    s.solve()
    o = Oracle(relax_type=relax, relax_order=0, max_levels=1, tolerance=1e-8, kspace=50, max_iterations=400)
    o.set_csr(A.indptr, A.indices, A.data, b)
    o.setup()
    o.solve()
    r = np.random.default_rng(4).standard_normal(A.shape[0])
    z, zo = vcycle_gpu(b200, s, r), o.vcycle(r)
    assert np.abs(z - zo).max() / np.abs(zo).max() <= VAL_RTOL
    assert abs(s.solver.iterations - o.iterations) <= 1
    assert np.abs(s.solution() - 1.0).max() < 1e-5
    s.destroy_system()


def test_structurally_nonsymmetric_matrix(b200):
    """a_ij stored without a_ji: the out-of-place sweep must still read the OLD value of a later row"""
    import scipy.sparse as sp
    rng = np.random.default_rng(9)
    n = 400
    M = sp.random(n, n, density=0.02, random_state=rng, format="lil")
    M = -abs(M.tocsr())
    M.setdiag(0)
    M.eliminate_zeros()
    A = (M + sp.diags(np.asarray(abs(M).sum(axis=1)).ravel() + 1.0)).tocsr()
    A.sort_indices()
    b = A @ np.ones(n)
    for relax in (3, 6, 8):
        s = csr_system(b200, A, b, relax_type=relax, relax_order=0, max_levels=1)
        s.solve()
        o = Oracle(relax_type=relax, relax_order=0, max_levels=1, tolerance=1e-8, kspace=50, max_iterations=400)
        o.set_csr(A.indptr, A.indices, A.data, b)
        o.setup()
        o.solve()
        r = rng.standard_normal(n)
        z, zo = vcycle_gpu(b200, s, r), o.vcycle(r)
        assert np.abs(z - zo).max() / np.abs(zo).max() <= 1e-11
        assert abs(s.solver.iterations - o.iterations) <= 1
        s.destroy_system()


@pytest.mark.parametrize("n,grid,relax,order", [((8, 8, 8), (2, 1, 1), 8, 1), ((6, 6, 6), (2, 2, 1), 6, 0), ((6, 5, 4), (2, 2, 2), 3, 1)])
def test_multirank_hybrid_gauss_seidel(b200, n, grid, relax, order):
    """Gauss-Seidel inside a rank, Jacobi across ranks: N processes against the oracle with N virtual ranks"""
    nranks = grid[0] * grid[1] * grid[2]
    res = run_ranks(nranks, settings(n, grid, 7, strong_threshold=0.25, relax_type=relax, relax_order=order))
    o = Oracle(strong_threshold=0.25, relax_type=relax, relax_order=order)
    o.threads(8)
    o.gen_laplace(n, p=grid, stencil=7)
    o.setup()
    xo = o.solve()
    for l in range(o.num_levels):
        np.testing.assert_allclose(np.concatenate([d[f"sm{l}"] for d in res]), o.smoother(l), rtol=1e-13)
    x = np.concatenate([d["x"] for d in res])
    its = int(res[0]["iterations"])
    assert abs(its - o.iterations) <= 1, (its, o.iterations)
    assert float(res[0]["relres"]) <= max(o.relres * 1.5, 1e-8)
    if its == o.iterations:
        assert np.abs(x - xo).max() / np.abs(xo).max() <= 1e-8
    assert np.abs(x - 1.0).max() < 1e-5
