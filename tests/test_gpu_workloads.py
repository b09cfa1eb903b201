"""GPU parity tests on the two nalu-wind-like workloads of BASELINE.json configs[3] / configs[4] (SURVEY.md §8d inputs
4 and 5), scaled down so that the CPU oracle finishes in seconds: perturbed-hex 27-point pressure Poisson with the
Chebyshev smoother, and the 3-DOF nonsymmetric momentum system with a smoother-only (max_levels 1) preconditioner.
The systems enter through the IJ path as unsorted COO triplets, like the file loaders hand them over
(/root/reference/src/HypreSystem.cpp:897-1015)."""
import numpy as np
import pytest

from oracle_lib import Oracle
from workloads import momentum3, poisson27_hex

pytestmark = pytest.mark.gpu

VAL_RTOL = 1e-10  # north_star: operator values and solutions within 1e-10 relative


def make_system(b200, A, b, amg, kspace=50, tol=1e-8, orth="mgs"):
    settings = {
        "linear_system": {"type": "csr", "ptr": A.indptr, "col": A.indices, "val": A.data, "rhs": b},
        "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": tol, "kspace": kspace,
                            "max_iterations": 400, "orthogonalization": orth},
        "boomeramg_settings": amg,
    }
    s = b200.HypreSystem(settings)
    s.setup_precon_and_solver()
    s.load()
    return s


def make_oracle(A, b, **opts):
    o = Oracle(tolerance=1e-8, kspace=50, max_iterations=400, **opts)
    o.threads(8)
    o.set_csr(A.indptr, A.indices, A.data, b)
    return o


def same_csr(got, ref, what, values=True):
    assert np.array_equal(got[0], ref[0]), f"{what}: row pointers differ"
    assert np.array_equal(got[1], ref[1].astype(np.int64)), f"{what}: pattern / entry order differs"
    if values and ref[2] is not None and len(ref[2]):
        err = np.abs(got[2] - ref[2]).max() / np.abs(ref[2]).max()
        assert err <= VAL_RTOL, f"{what}: values differ by {err:.3e} relative"


def check_hierarchy(amg, o):
    assert amg.num_levels == o.num_levels
    for l in range(o.num_levels):
        same_csr(amg.level_matrix(l, "A"), o.matrix(l, "A"), f"A[{l}]")
        if l + 1 < o.num_levels:
            assert np.array_equal(amg.level_cf(l), o.cf(l)), f"CF[{l}] differs"
            same_csr(amg.level_matrix(l, "S"), o.matrix(l, "S"), f"S[{l}]", values=False)
            same_csr(amg.level_matrix(l, "P"), o.matrix(l, "P"), f"P[{l}]")
            same_csr(amg.level_matrix(l, "R"), o.matrix(l, "R"), f"R[{l}]")


def check_solve(s, o, xo):
    x = s.solution()
    assert s.solver.converged == 1
    assert abs(s.solver.iterations - o.iterations) <= 1, (s.solver.iterations, o.iterations)
    assert s.solver.relres <= max(o.relres * (1 + 1e-6), 1e-8)
    assert np.abs(x - 1.0).max() < 1e-5  # rhs = A 1
    if s.solver.iterations == o.iterations:
        assert np.abs(x - xo).max() / np.abs(xo).max() <= 1e-8


@pytest.mark.parametrize("n,theta", [(12, 0.25), (14, 0.5)])
def test_poisson27_hex_hierarchy_bit_exact(b200, n, theta):
    """mixed-sign off-diagonals (the trilinear-hex Laplacian has positive face couplings once the mesh is perturbed)"""
    A, b = poisson27_hex(n)
    s = make_system(b200, A, b, dict(strong_threshold=theta))
    s.solve()
    o = make_oracle(A, b, strong_threshold=theta)
    o.setup()
    check_hierarchy(s.precond, o)
    s.destroy_system()


@pytest.mark.parametrize("relax", [16, 18])
def test_poisson27_hex_solve_parity(b200, relax):
    A, b = poisson27_hex(14)
    s = make_system(b200, A, b, dict(relax_type=relax))
    s.solve()
    o = make_oracle(A, b, relax_type=relax)
    o.setup()
    xo = o.solve()
    if relax == 16:
        for l in range(o.num_levels - 1):
            np.testing.assert_allclose(s.precond.level_cheby(l)[:2], o.cheby(l)[:2], rtol=1e-9)
    check_solve(s, o, xo)
    s.destroy_system()


def test_momentum3_hierarchy_bit_exact(b200):
    """nonsymmetric operator with 3x3 node blocks through the full setup (strength, PMIS, ext+i, RAP)"""
    A, b = momentum3(8)
    s = make_system(b200, A, b, dict(strong_threshold=0.25))
    s.solve()
    o = make_oracle(A, b, strong_threshold=0.25)
    o.setup()
    check_hierarchy(s.precond, o)
    xo = o.solve()
    check_solve(s, o, xo)
    s.destroy_system()


@pytest.mark.parametrize("relax", [7, 18])
def test_momentum3_smoother_only_preconditioner(b200, relax):
    """BASELINE configs[4]: 'Jacobi preconditioner' = BoomerAMG with max_levels 1"""
    A, b = momentum3(10)
    s = make_system(b200, A, b, dict(relax_type=relax, max_levels=1))
    s.solve()
    assert s.precond.num_levels == 1
    o = make_oracle(A, b, relax_type=relax, max_levels=1)
    o.setup()
    xo = o.solve()
    check_solve(s, o, xo)
    s.destroy_system()


def test_momentum3_cgs2_krylov_dominated(b200):
    """the fused CGS2 Arnoldi step on a long restart cycle (kspace 50, ~30 iterations) against MGS"""
    A, b = momentum3(10)
    its = {}
    for orth in ("mgs", "cgs2"):
        s = make_system(b200, A, b, dict(relax_type=18, max_levels=1), orth=orth)
        s.solve()
        its[orth] = s.solver.iterations
        assert s.solver.converged == 1
        assert np.abs(s.solution() - 1.0).max() < 1e-5
        s.destroy_system()
    assert abs(its["mgs"] - its["cgs2"]) <= 1


@pytest.mark.parametrize("kind,n", [("poisson27_hex", 9), ("momentum3", 8)])
def test_device_generated_workload_equals_numpy_builder(b200, kind, n):
    """BASELINE.json configs[3] / [4] at their stated size are generated on the device (csrc/workload_rows.cuh); at small size
    the device-generated system, assembled through the IJ path, is the matrix of the numpy builder (momentum: every bit;
    hex Poisson: up to the summation order of the element contributions), and it solves to x == 1"""
    import scipy.sparse as sp
    A, b = poisson27_hex(n) if kind == "poisson27_hex" else momentum3(n)
    amg = {"relax_type": 16, "strong_threshold": 0.5} if kind == "poisson27_hex" else {"max_levels": 1, "relax_type": 18}
    s = b200.HypreSystem({"linear_system": {"type": kind, "n": n},
                          "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50,
                                              "max_iterations": 400},
                          "boomeramg_settings": amg})
    s.setup_precon_and_solver()
    s.load()
    s.assemble_system()
    ptr, col, val, _ = b200.parcsr_to_host(s.A.object())
    G = sp.csr_matrix((val, col, ptr), shape=A.shape)
    G.sort_indices()
    R = A.tocsr()
    R.sort_indices()
    assert np.array_equal(G.indptr, R.indptr) and np.array_equal(G.indices, R.indices)
    scale = np.abs(R.data).max()
    if kind == "momentum3":
        assert np.array_equal(G.data, R.data) and np.array_equal(s.b.get_values(), b)
    else:
        assert np.abs(G.data - R.data).max() <= 1e-13 * scale and np.abs(s.b.get_values() - b).max() <= 1e-12 * scale
    s.solve()
    assert s.solver.converged == 1 and np.abs(s.solution() - 1.0).max() < 1e-5
    s.destroy_system()
