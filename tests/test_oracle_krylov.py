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


# ---- ILU(0) (hypre ILU type 0, fill 0: block-Jacobi ILU(0); HypreSystem.cpp:328-370,457-497) ---------------------------------
def ilu_lu(o, lo=0, hi=None):
    """L (unit lower) and U of one diagonal block as scipy matrices, from the oracle's factor values"""
    ptr, col, _, shape = o.matrix(0, "A")
    val, dinv = o.ilu_factors()
    hi = shape[0] if hi is None else hi
    rows = np.repeat(np.arange(shape[0]), np.diff(ptr))
    blk = (rows >= lo) & (rows < hi) & (col >= lo) & (col < hi)
    F = sp.csr_matrix((val[blk], (rows[blk] - lo, col[blk] - lo)), shape=(hi - lo, hi - lo))
    Lm = sp.tril(F, -1) + sp.identity(hi - lo)
    Um = sp.triu(F, 0)
    return Lm.tocsr(), Um.tocsr(), dinv[lo:hi]


@pytest.mark.parametrize("which", ["lap7", "lap27", "advdiff"])
def test_ilu0_defining_property(which):
    """ILU(0): (L U)_ij = a_ij on the pattern of A; 1 / u_ii stored as the inverted pivot"""
    if which == "advdiff":
        A, b = advection_diffusion(7)
        o = Oracle()
        o.set_csr(A.indptr, A.indices, A.data, b)
    else:
        o = system(7, 7 if which == "lap7" else 27)
        A = o.scipy(0, "A")
    o.ilu_setup()
    Lm, Um, dinv = ilu_lu(o)
    np.testing.assert_allclose(dinv, 1.0 / Um.diagonal(), rtol=1e-15)
    P = (Lm @ Um).tocsr()
    Ac = A.tocoo()
    got = np.asarray(P[Ac.row, Ac.col]).ravel()
    np.testing.assert_allclose(got, Ac.data, rtol=1e-12, atol=1e-13)
    # the application is the two triangular solves
    rng = np.random.default_rng(4)
    r = rng.standard_normal(A.shape[0])
    z = o.ilu_apply(r)
    zref = spla.spsolve_triangular(Um, spla.spsolve_triangular(Lm, r, lower=True), lower=False)
    np.testing.assert_allclose(z, zref, rtol=1e-11, atol=1e-13)


def test_ilu0_is_exact_lu_for_a_tridiagonal_matrix():
    n = 50
    A = sp.diags([-np.ones(n - 1), 2.5 * np.ones(n), -1.5 * np.ones(n - 1)], [-1, 0, 1]).tocsr()
    b = A @ np.ones(n)
    o = Oracle(tolerance=1e-12, max_iterations=5)
    o.set_csr(A.indptr, A.indices, A.data, b)
    o.ilu_setup()
    x = o.solve(method="ilu")
    assert o.iterations == 1 and np.abs(x - 1.0).max() < 1e-12


def test_ilu0_block_jacobi_on_virtual_ranks():
    """every rank factors its own diagonal block; couplings to other ranks are ignored"""
    o = Oracle()
    o.gen_laplace((5, 5, 5), p=(2, 1, 1), stencil=7)
    o.ilu_setup()
    A = o.scipy(0, "A")
    for lo, hi in ((0, 125), (125, 250)):
        Lm, Um, _ = ilu_lu(o, lo, hi)
        B = A[lo:hi][:, lo:hi].tocoo()
        got = np.asarray((Lm @ Um).tocsr()[B.row, B.col]).ravel()
        np.testing.assert_allclose(got, B.data, rtol=1e-12, atol=1e-13)


@pytest.mark.parametrize("method", ["gmres", "cg", "bicg", "fgmres", "ilu"])
def test_ilu_preconditioned_methods_converge(method):
    tol = 1e-8
    o = system(10, 7, tolerance=tol, kspace=30, max_iterations=300)
    A, b = o.scipy(0, "A"), o.rhs()
    o.ilu_setup()
    x = o.solve(method=method)
    assert o.L.ora_converged(o.h) == 1
    assert np.linalg.norm(b - A @ x) / np.linalg.norm(b) < (50 * tol if method == "cg" else 1.0001 * tol)
    if method != "ilu":   # ILU(0) preconditioning beats no preconditioning
        p = system(10, 7, tolerance=tol, kspace=30, max_iterations=300, preconditioner=0)
        p.solve(method=method)
        assert o.iterations < p.iterations
