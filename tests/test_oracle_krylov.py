"""CPU tests that pin the oracle's sibling Krylov methods (oracle/krylov.inc: PCG, BiCGSTAB, FlexGMRES, COGMRES,
BoomerAMG as solver -- the `method` values of /root/reference/src/HypreSystem.cpp:69-85).  hypre's sources are not in
/root/reference, so the pins are independent computations of the same algorithms: scipy's cg / bicgstab iterates, a
direct solve, and the algebraic identities that relate the GMRES variants to each other."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from oracle_lib import Oracle

METHODS = ["gmres", "fgmres", "cogmres", "cg", "bicg", "boomeramg"]


def system(n=10, stencil=7, **opts):
    o = Oracle(**opts)
    o.threads(4)
    o.gen_laplace(n, stencil=stencil)
    return o


def advection_diffusion(n=8, seed=5):
    """small nonsymmetric, diagonally dominant matrix (upwinded advection-diffusion, the shape of BASELINE config 5)"""
    rng = np.random.default_rng(seed)
    N = n ** 3
    idx = np.arange(N).reshape(n, n, n)
    rows, cols, vals = [], [], []
    for axis, vel in enumerate((1.0, 0.5, 0.25)):
        lo = np.take(idx, range(0, n - 1), axis=axis).ravel()
        hi = np.take(idx, range(1, n), axis=axis).ravel()
        rows += [hi, lo]; cols += [lo, hi]
        vals += [np.full(lo.size, -1.0 - vel), np.full(lo.size, -1.0)]     # upwind: the inflow neighbour carries the advection
    A = sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(N, N))
    A = A + sp.diags(-np.asarray(A.sum(axis=1)).ravel() + 0.05 + 0.01 * rng.random(N))
    A = A.tocsr()
    A.sort_indices()
    return A, A @ np.ones(N)


@pytest.mark.parametrize("method", METHODS)
@pytest.mark.parametrize("precond", [1, 0])
def test_every_method_solves_the_generator_system(method, precond):
    if method == "boomeramg" and not precond:
        pytest.skip("BoomerAMG as solver needs the hierarchy")
    tol = 1e-9
    o = system(10, 7, tolerance=tol, kspace=25, max_iterations=400)
    if precond:
        o.setup()
    x = o.solve(method=method)
    A, b = o.scipy(0, "A"), o.rhs()
    assert o.L.ora_converged(o.h) == 1
    true_rel = np.linalg.norm(b - A @ x) / np.linalg.norm(b)
    if method == "cg":      # PCG stops on the preconditioner energy norm <C r, r> / <C b, b> (hypre default two_norm = 0)
        assert o.relres < tol and true_rel < 50 * tol
    else:                   # the others confirm ||b - A x|| <= tol ||b|| with the true residual before stopping
        assert true_rel <= tol * 1.0001
        assert abs(o.relres - true_rel) <= 1e-3 * tol
    assert np.abs(x - 1.0).max() < 1e-6
    assert len(o.history()) == o.iterations + 1


def test_pcg_unpreconditioned_iterates_match_scipy_cg():
    o = system(7, 7, tolerance=1e-30, preconditioner=0)
    A, b = o.scipy(0, "A"), o.rhs()
    its = []
    spla.cg(A, b, x0=np.zeros_like(b), rtol=1e-30, maxiter=12, callback=lambda xk: its.append(xk.copy()))
    for k in (1, 2, 5, 12):
        o.set(max_iterations=k)
        xk = o.solve(method="cg")
        assert o.iterations == k
        np.testing.assert_allclose(xk, its[k - 1], rtol=1e-10, atol=1e-12)
    # without a preconditioner the PCG norm is the 2-norm: history = ||r_k||
    o.set(max_iterations=5)
    x5 = o.solve(method="cg")
    assert abs(o.history()[-1] - np.linalg.norm(b - A @ x5)) <= 1e-10 * np.linalg.norm(b)


def test_bicgstab_unpreconditioned_iterates_match_scipy_bicgstab():
    A, b = advection_diffusion(7)
    o = Oracle(tolerance=1e-30, preconditioner=0)
    o.set_csr(A.indptr, A.indices, A.data, b)
    its = []
    spla.bicgstab(A, b, x0=np.zeros_like(b), rtol=1e-30, maxiter=10, callback=lambda xk: its.append(xk.copy()))
    for k in (1, 3, 10):
        o.set(max_iterations=k)
        xk = o.solve(method="bicg")
        np.testing.assert_allclose(xk, its[k - 1], rtol=1e-8, atol=1e-11)
        assert abs(o.history()[-1] - np.linalg.norm(b - A @ xk)) <= 1e-9 * np.linalg.norm(b)


def test_bicgstab_and_gmres_family_on_a_nonsymmetric_matrix():
    A, b = advection_diffusion(8)
    xd = spla.spsolve(A.tocsc(), b)
    for method in ("bicg", "gmres", "fgmres", "cogmres"):
        o = Oracle(tolerance=1e-10, kspace=40, max_iterations=300, preconditioner=0)
        o.set_csr(A.indptr, A.indices, A.data, b)
        x = o.solve(method=method)
        assert o.L.ora_converged(o.h) == 1, method
        assert np.linalg.norm(b - A @ x) <= 1e-10 * np.linalg.norm(b) * 1.0001, method
        assert np.abs(x - xd).max() < 1e-8, method


@pytest.mark.parametrize("precond", [1, 0])
def test_flexgmres_and_cogmres_equal_gmres_for_a_fixed_preconditioner(precond):
    """With a fixed (linear) preconditioner FlexGMRES builds the same Krylov space as right-preconditioned GMRES, and
    classical Gram-Schmidt equals modified Gram-Schmidt in exact arithmetic: residual histories agree to rounding
    while the basis is well conditioned, and the solutions agree to the tolerance."""
    ref = None
    for method, cgs in (("gmres", 0), ("fgmres", 0), ("cogmres", 0), ("cogmres", 2)):
        o = system(9, 7, tolerance=1e-10, kspace=12, max_iterations=200, cgs=cgs)   # kspace 12 forces restarts
        if precond:
            o.setup()
        else:
            o.set(preconditioner=0)
        x = o.solve(method=method)
        h = o.history()
        if ref is None:
            ref = (x, h, o.iterations)
            continue
        assert abs(o.iterations - ref[2]) <= 1, (method, cgs)
        k = min(len(h), len(ref[1]), 13)               # the first restart cycle
        np.testing.assert_allclose(h[:k], ref[1][:k], rtol=1e-7)
        assert np.abs(x - ref[0]).max() < 1e-7


def test_amg_solver_is_the_stationary_vcycle_iteration():
    o = system(9, 7, tolerance=1e-8, max_iterations=60)
    o.setup()
    x = o.solve(method="boomeramg")
    A, b = o.scipy(0, "A"), o.rhs()
    y = np.zeros_like(b)
    for _ in range(o.iterations):
        y = y + o.vcycle(b - A @ y)
    np.testing.assert_allclose(x, y, rtol=1e-12, atol=1e-13)
    h = o.history()
    assert np.all(h[1:] < h[:-1])                                       # a convergent V-cycle: monotone residuals
    assert h[-1] / np.linalg.norm(b) < 1e-8 <= h[-2] / np.linalg.norm(b)  # stops at the first cycle below tol


def test_pcg_zero_rhs_and_max_iter():
    o = system(5, 7, tolerance=1e-8, preconditioner=0)
    A = o.scipy(0, "A")
    z = Oracle(tolerance=1e-8, preconditioner=0)
    z.set_csr(A.indptr, A.indices, A.data, np.zeros(A.shape[0]))
    x = z.solve(x0=np.ones(A.shape[0]), method="cg")
    assert np.all(x == 0.0) and z.iterations == 0                        # hypre_PCGSolve: b == 0 => x = 0
    o.set(max_iterations=3)
    o.solve(method="cg")
    assert o.iterations == 3 and o.L.ora_converged(o.h) == 0
