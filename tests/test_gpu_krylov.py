"""GPU parity tests of the sibling Krylov methods: FlexGMRES, COGMRES, PCG, BiCGSTAB and BoomerAMG as
solver through the C ABI (HYPRE_ParCSRFlexGMRES* / COGMRES* / PCG* / BiCGSTAB* / HYPRE_BoomerAMGSolve) against the CPU
oracle (oracle/krylov.inc) on the same inputs.  Bars as in test_gpu_parity.py: iterations within +-1, residual history
and solution to rounding when the counts agree, final relative residual at or below the oracle's."""
import numpy as np
import pytest

from oracle_lib import Oracle
from test_oracle_krylov import advection_diffusion

pytestmark = pytest.mark.gpu

TOL = 1e-8


def gpu_system(b200, method, precond, linear_system, kspace=30, max_it=400, tol=TOL, **extra):
    settings = {
        "linear_system": linear_system,
        "solver_settings": dict({"method": method, "preconditioner": "boomeramg" if precond else "none", "tolerance": tol,
                                 "kspace": kspace, "max_iterations": max_it}, **extra),
        "boomeramg_settings": {},
    }
    s = b200.HypreSystem(settings)
    s.setup_precon_and_solver()
    s.load()
    return s


def laplace_system(n, stencil=7):
    return {"type": "build_27pt_stencil" if stencil == 27 else "build_7pt_stencil", "nx": n, "ny": n, "nz": n}


def compare(s, o, xo, exact_history=True):
    x = s.solution()
    assert s.solver.converged == 1
    assert abs(s.solver.iterations - o.iterations) <= 1, (s.solver.iterations, o.iterations)
    assert s.solver.relres <= max(o.relres * (1 + 1e-6), TOL)
    if s.solver.iterations == o.iterations:
        assert np.abs(x - xo).max() / np.abs(xo).max() <= 1e-8
        if exact_history:
            hg, ho = s.solver.history(), o.history()
            assert len(hg) == len(ho)
            big = ho > 1e-3 * TOL * ho[0]
            np.testing.assert_allclose(hg[big], ho[big], rtol=1e-5)
    return x


@pytest.mark.parametrize("method", ["fgmres", "cogmres", "cg", "bicg"])
@pytest.mark.parametrize("precond", [1, 0])
def test_method_parity_laplace(b200, method, precond):
    n = 20 if precond else 14
    s = gpu_system(b200, method, precond, laplace_system(n), kspace=20)
    s.solve()
    o = Oracle(tolerance=TOL, kspace=20, max_iterations=400, preconditioner=precond)
    o.threads(8)
    o.gen_laplace(n, stencil=7)
    if precond:
        o.setup()
    xo = o.solve(method=method)
    # BiCGSTAB's recurrences amplify rounding differences (tree vs blocked sums): compare counts and solutions only
    x = compare(s, o, xo, exact_history=(method != "bicg"))
    assert np.abs(x - 1.0).max() < 1e-5
    s.destroy_system()


@pytest.mark.parametrize("cgs", [0, 1, 2])
def test_cogmres_cgs_variants(b200, cgs):
    n = 16
    s = gpu_system(b200, "cogmres", 1, laplace_system(n), kspace=8, cgs=cgs)   # kspace 8 forces restarts
    s.solve()
    o = Oracle(tolerance=TOL, kspace=8, max_iterations=400, cgs=cgs)
    o.gen_laplace(n, stencil=7)
    o.setup()
    xo = o.solve(method="cogmres")
    compare(s, o, xo)
    s.destroy_system()


def test_flexgmres_restarts(b200):
    n = 16
    s = gpu_system(b200, "fgmres", 1, laplace_system(n, 27), kspace=5)
    s.solve()
    o = Oracle(tolerance=TOL, kspace=5, max_iterations=400)
    o.gen_laplace(n, stencil=27)
    o.setup()
    xo = o.solve(method="fgmres")
    assert o.iterations > 5                                                    # at least one restart happened
    compare(s, o, xo)
    s.destroy_system()


@pytest.mark.parametrize("method", ["bicg", "fgmres", "cogmres"])
def test_nonsymmetric_matrix(b200, method):
    A, b = advection_diffusion(10)
    ls = {"type": "csr", "ptr": A.indptr, "col": A.indices, "val": A.data, "rhs": b}
    s = gpu_system(b200, method, 0, ls, kspace=40)
    s.solve()
    o = Oracle(tolerance=TOL, kspace=40, max_iterations=400, preconditioner=0)
    o.set_csr(A.indptr, A.indices, A.data, b)
    xo = o.solve(method=method)
    x = compare(s, o, xo, exact_history=(method != "bicg"))
    assert np.linalg.norm(b - A @ x) <= TOL * np.linalg.norm(b) * 1.0001
    s.destroy_system()


@pytest.mark.parametrize("relax,order", [(18, 0), (6, 1)])
def test_boomeramg_as_solver(b200, relax, order):
    n = 16
    s = gpu_system(b200, "boomeramg", 1, laplace_system(n), max_it=100, relax_type=relax, relax_order=order, strong_threshold=0.25)
    s.solve()
    o = Oracle(tolerance=TOL, max_iterations=100, relax_type=relax, relax_order=order, strong_threshold=0.25)
    o.gen_laplace(n, stencil=7)
    o.setup()
    xo = o.solve(method="boomeramg")
    x = s.solution()
    assert s.solver.iterations == o.iterations
    assert s.solver.relres < TOL and abs(s.solver.relres - o.relres) <= 1e-6 * o.relres + 1e-14
    assert np.abs(x - xo).max() / np.abs(xo).max() <= 1e-9
    s.destroy_system()


def test_pcg_two_norm_and_max_iter(b200):
    n = 12
    s = gpu_system(b200, "cg", 0, laplace_system(n), max_it=5)
    s.solve()
    assert s.solver.iterations == 5 and s.solver.converged == 0                # HYPRE_ERROR_CONV reported, not fatal
    o = Oracle(tolerance=TOL, max_iterations=5, preconditioner=0)
    o.gen_laplace(n, stencil=7)
    xo = o.solve(method="cg")
    np.testing.assert_allclose(s.solution(), xo, rtol=1e-10, atol=1e-12)       # same 5 CG iterates
    s.destroy_system()
    # two_norm = 1: the stopping test uses ||r||_2 / ||b||_2
    s = gpu_system(b200, "cg", 1, laplace_system(16))
    assert b200.lib().HYPRE_ParCSRPCGSetTwoNorm(s.solver.h, 1) == 0
    s.solve()
    x = s.solution()
    o = Oracle()
    o.gen_laplace(16, stencil=7)
    A, b = o.scipy(0, "A"), o.rhs()
    true_rel = np.linalg.norm(b - A @ x) / np.linalg.norm(b)
    assert s.solver.converged == 1 and true_rel < TOL * 1.001 and abs(true_rel - s.solver.relres) < 1e-3 * TOL
    s.destroy_system()


@pytest.mark.parametrize("method", ["cg", "bicg", "fgmres"])
def test_larger_size_properties(b200, method):
    """96^3 rows: x == 1 (generator known answer) and an independent residual through the public matvec"""
    n = 96
    s = gpu_system(b200, method, 1, laplace_system(n), kspace=30)
    s.solve()
    x = s.solution()
    assert s.solver.converged == 1
    assert np.abs(x - 1.0).max() < 1e-5
    r = b200.IJVector(0, n ** 3 - 1)
    b200.matvec(s.A, s.x, r)
    bb = s.b.get_values()
    rel = np.linalg.norm(bb - r.get_values()) / np.linalg.norm(bb)
    assert rel <= (50 * TOL if method == "cg" else 1.01 * TOL)                # PCG stops on the energy norm
    r.destroy()
    s.destroy_system()


@pytest.mark.parametrize("method", ["gmres", "fgmres", "cogmres", "cg", "bicg", "boomeramg"])
def test_zero_rhs_and_exact_initial_guess(b200, method):
    """edge cases of the stopping logic: b = 0 gives x = 0 without iterating; an exact initial guess is accepted as is"""
    n = 10
    s = gpu_system(b200, method, 1, laplace_system(n))
    s.b.set_values(np.zeros(n ** 3))
    s.solve()
    assert s.solver.iterations == 0
    assert np.all(s.solution() == 0.0)
    s.destroy_system()
    s = gpu_system(b200, method, 1, laplace_system(n))
    s.x.set_values(np.ones(n ** 3))            # rhs = A 1 (generator): the initial guess is the solution
    s.solve()
    # hypre_PCGSolve has no test before its first step: it enters iteration 1, finds <A p, p> = 0 and stops there
    assert s.solver.iterations == (1 if method == "cg" else 0)
    np.testing.assert_allclose(s.solution(), 1.0, rtol=0, atol=1e-14)
    s.destroy_system()


@pytest.mark.parametrize("method", ["gmres", "fgmres", "cogmres", "bicg"])
def test_iteration_limit_is_reported_not_fatal(b200, method):
    """max_iterations reached: HYPRE_ERROR_CONV semantics (flag raised, iterate returned), the mini-app ignores the code"""
    n = 12
    s = gpu_system(b200, method, 0, laplace_system(n), kspace=4, max_it=6)
    s.solve()
    assert s.solver.iterations == 6 and s.solver.converged == 0
    o = Oracle(tolerance=TOL, kspace=4, max_iterations=6, preconditioner=0)
    o.gen_laplace(n, stencil=7)
    xo = o.solve(method=method)
    assert o.iterations == 6
    np.testing.assert_allclose(s.solution(), xo, rtol=1e-8, atol=1e-10)
    s.destroy_system()
