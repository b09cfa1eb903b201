"""GPU parity tests of the ILU solver / preconditioner (/root/reference/src/HypreSystem.cpp:328-370,
457-497 with the mini-app's defaults = block-Jacobi ILU(0)) through HYPRE_ILU* of the C ABI against the CPU oracle
(oracle/krylov.inc: ilu0_setup / ilu_apply / ilu_solver).  Bars: factor values within 1e-12 relative (same elimination
order, unfused arithmetic), preconditioned iteration counts within +-1, solutions within 1e-8, residual at or below the
oracle's."""
import numpy as np
import pytest

from oracle_lib import Oracle
from test_oracle_krylov import advection_diffusion
from workloads import momentum3

pytestmark = pytest.mark.gpu

TOL = 1e-8


def gpu_system(b200, linear_system, method="gmres", precond="ilu", kspace=40, max_it=400, tol=TOL, ilu=None, **extra):
    settings = {
        "linear_system": linear_system,
        "solver_settings": dict({"method": method, "preconditioner": precond, "tolerance": tol, "kspace": kspace,
                                 "max_iterations": max_it}, **extra),
        "ilu_preconditioner_settings": ilu or {},
    }
    s = b200.HypreSystem(settings)
    s.setup_precon_and_solver()
    s.load()
    return s


def csr_system(A, b):
    return {"type": "csr", "ptr": A.indptr, "col": A.indices, "val": A.data, "rhs": b}


def laplace(n, stencil):
    o = Oracle()
    o.gen_laplace(n, stencil=stencil)
    return o.scipy(0, "A"), o.rhs()


def matrices():
    yield "lap7", laplace(14, 7)
    yield "lap27", laplace(9, 27)
    yield "advdiff", advection_diffusion(9)
    yield "momentum3", momentum3(6)


@pytest.mark.parametrize("name,Ab", list(matrices()), ids=[m[0] for m in matrices()])
def test_ilu0_factors_match_oracle(b200, name, Ab):
    A, b = Ab
    s = gpu_system(b200, csr_system(A, b))
    s.assemble_system()
    s.precond.setup(s.A)
    o = Oracle()
    o.set_csr(A.indptr, A.indices, A.data, b)
    o.ilu_setup()
    vo, do = o.ilu_factors()
    vg, dg = s.precond.factors(len(vo), A.shape[0])   # one rank: A_diag has the pattern / order of the oracle's matrix
    assert np.abs(vg - vo).max() <= 1e-12 * np.abs(vo).max()
    assert np.abs(dg - do).max() <= 1e-12 * np.abs(do).max()
    s.destroy_system()


@pytest.mark.parametrize("method", ["gmres", "fgmres", "cg", "bicg"])
def test_ilu_preconditioned_solve_parity(b200, method):
    n = 16
    s = gpu_system(b200, {"type": "build_7pt_stencil", "nx": n, "ny": n, "nz": n}, method=method, orthogonalization="mgs")
    s.solve()
    o = Oracle(tolerance=TOL, kspace=40, max_iterations=400)
    o.gen_laplace(n, stencil=7)
    o.ilu_setup()
    xo = o.solve(method=method)
    x = s.solution()
    assert s.solver.converged == 1
    assert abs(s.solver.iterations - o.iterations) <= 1, (s.solver.iterations, o.iterations)
    assert s.solver.relres <= max(o.relres * (1 + 1e-6), TOL)
    assert np.abs(x - 1.0).max() < 1e-5
    if s.solver.iterations == o.iterations:
        assert np.abs(x - xo).max() / np.abs(xo).max() <= 1e-8
    s.destroy_system()


def test_ilu_on_the_nonsymmetric_momentum_system(b200):
    A, b = momentum3(8)
    s = gpu_system(b200, csr_system(A, b), method="bicg")
    s.solve()
    o = Oracle(tolerance=TOL, max_iterations=400)
    o.set_csr(A.indptr, A.indices, A.data, b)
    o.ilu_setup()
    xo = o.solve(method="bicg")
    x = s.solution()
    assert s.solver.converged == 1 and abs(s.solver.iterations - o.iterations) <= 1
    assert np.linalg.norm(b - A @ x) <= TOL * np.linalg.norm(b) * 1.0001
    assert np.abs(x - xo).max() / np.abs(xo).max() <= 1e-7
    s.destroy_system()


@pytest.mark.parametrize("tol,max_it", [(1e-4, 400), (0.0, 7)])
def test_ilu_as_solver(b200, tol, max_it):
    """`method: ilu`: x += (LU)^{-1} (b - A x) until the relative residual drops below tol, or exactly max_iterations sweeps"""
    n = 10
    s = gpu_system(b200, {"type": "build_7pt_stencil", "nx": n, "ny": n, "nz": n}, method="ilu", precond="none", max_it=max_it, tol=tol)
    s.solve()
    o = Oracle(tolerance=tol, max_iterations=max_it)
    o.gen_laplace(n, stencil=7)
    o.ilu_setup()
    xo = o.solve(method="ilu")
    assert s.solver.iterations == o.iterations
    if tol > 0:
        assert s.solver.relres < tol and abs(s.solver.relres - o.relres) <= 1e-8 * o.relres
    np.testing.assert_allclose(s.solution(), xo, rtol=1e-10, atol=1e-12)
    s.destroy_system()


def test_unsupported_ilu_variants_are_rejected(b200):
    n = 6
    for bad in ({"ilu_type": 10}, {"fill": 1}, {"local_reordering": 1}, {"trisolve": 0}):
        s = gpu_system(b200, {"type": "build_7pt_stencil", "nx": n, "ny": n, "nz": n}, ilu=bad)
        with pytest.raises(b200.HypreError):
            s.solve()
        b200.lib().HYPRE_ClearAllErrors()
        s.destroy_system()


def test_larger_size_properties(b200):
    """64^3 rows: GMRES + ILU(0) reaches the tolerance, x == 1, fewer iterations than without a preconditioner"""
    n = 64
    its = {}
    for precond in ("ilu", "none"):
        s = gpu_system(b200, {"type": "build_7pt_stencil", "nx": n, "ny": n, "nz": n}, precond=precond, kspace=60, max_it=600)
        s.solve()
        assert s.solver.converged == 1 and np.abs(s.solution() - 1.0).max() < 1e-5
        its[precond] = s.solver.iterations
        s.destroy_system()
    assert its["ilu"] < 0.6 * its["none"]
