"""N-rank parity on ONE GPU: N processes share cuda:0 over the host-staged shm transport (NCCL refuses duplicate GPUs),
run the distributed path of the library (halo exchange, distributed PMIS / ext+i / RAP, replicated coarse solve, Krylov
all-reduces) and are compared with the oracle run with N virtual ranks (per-rank PMIS seeds 2747 + rank).

Bars (north_star): hierarchy bit-exact (CF, S / P / R / A_c patterns incl. entry order, level sizes), values and solution
within 1e-10 relative, iterations within +-1, final relative residual <= the oracle's (last-digit noise allowed)."""
import json
import os
import subprocess
import sys
import tempfile
import uuid

import numpy as np
import pytest

from oracle_lib import Oracle

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
VAL_RTOL = 1e-10


def run_ranks_nccl(nranks, cfg, timeout=600):
    """one rank per GPU under torchrun, NCCL transport (needs >= nranks GPUs)"""
    port = 29000 + (os.getpid() % 500)
    with tempfile.TemporaryDirectory() as tmp:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nranks}", "--master-addr", "127.0.0.1",
               "--master-port", str(port), os.path.join(HERE, "mr_worker.py"), "-1", str(nranks), "nccl", tmp, json.dumps(cfg)]
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
        assert p.returncode == 0, p.stdout[-4000:]
        return [dict(np.load(os.path.join(tmp, f"rank{r}.npz"))) for r in range(nranks)]


def n_gpus():
    import torch
    return torch.cuda.device_count()


def run_ranks(nranks, cfg, timeout=600):
    name = "/b200hypre-" + uuid.uuid4().hex[:12]
    with tempfile.TemporaryDirectory() as tmp:
        env = dict(os.environ, B200_SHM_TIMEOUT_S=os.environ.get("B200_SHM_TIMEOUT_S", "120"))   # a desynchronised rank fails in 2 min, not 5
        procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "mr_worker.py"), str(r), str(nranks), name, tmp, json.dumps(cfg)],
                                  stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env) for r in range(nranks)]
        outs = []
        failed = False
        for p in procs:
            try:
                o, _ = p.communicate(timeout=timeout)
            except subprocess.TimeoutExpired:
                failed = True
                for q in procs:
                    q.kill()
                o = "timeout"
            outs.append(o)
            failed |= p.returncode != 0
        assert not failed, "a rank failed:\n" + "\n-----\n".join(x[-3000:] for x in outs)
        return [dict(np.load(os.path.join(tmp, f"rank{r}.npz"))) for r in range(nranks)]


def stack_csr(res, key):
    """concatenate the per-rank row blocks (rank order = global row order)"""
    ptrs, cols, vals = [], [], []
    off = 0
    for d in res:
        p = d[f"{key}_ptr"].astype(np.int64)
        ptrs.append(p[:-1] + off)
        off += p[-1]
        cols.append(d[f"{key}_col"])
        if f"{key}_val" in d:
            vals.append(d[f"{key}_val"])
    ptr = np.concatenate(ptrs + [np.array([off])])
    return ptr, np.concatenate(cols), (np.concatenate(vals) if vals else None)


def check_csr(got, ref, what, values=True):
    gptr, gcol, gval = got
    rptr, rcol, rval = ref[0], ref[1], ref[2]
    assert np.array_equal(gptr, rptr.astype(np.int64)), f"{what}: row pointers differ"
    assert np.array_equal(gcol, rcol.astype(np.int64)), f"{what}: column pattern / entry order differs"
    if values and rval is not None and len(rval):
        err = np.abs(gval - rval).max() / np.abs(rval).max()
        assert err <= VAL_RTOL, f"{what}: values differ by {err:.3e} relative"


def settings(n, grid, stencil, **amg):
    return {
        "linear_system": {"type": "build_27pt_stencil" if stencil == 27 else "build_7pt_stencil", "nx": n[0], "ny": n[1], "nz": n[2],
                          "grid": list(grid)},
        "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 200},
        "boomeramg_settings": amg,
    }


CASES = [
    ((8, 8, 8), (2, 1, 1), 7, 0.25),
    ((6, 6, 6), (2, 2, 1), 7, 0.25),
    ((6, 5, 4), (2, 2, 2), 7, 0.25),
    ((6, 6, 6), (2, 1, 1), 27, 0.25),
    ((5, 5, 5), (2, 2, 2), 27, 0.57),
    ((12, 12, 12), (3, 1, 1), 7, 0.25),
    ((24, 24, 24), (2, 2, 1), 7, 0.25),
]


@pytest.mark.parametrize("n,grid,stencil,theta", CASES)
def test_multirank_hierarchy_and_solve(b200, n, grid, stencil, theta):
    nranks = grid[0] * grid[1] * grid[2]
    res = run_ranks(nranks, settings(n, grid, stencil, strong_threshold=theta))
    compare_with_oracle(res, n, grid, stencil, theta)


@pytest.mark.parametrize("n,grid,stencil,theta", [((16, 16, 16), (2, 1, 1), 7, 0.25), ((32, 32, 32), (2, 1, 1), 7, 0.25),
                                                 ((16, 16, 16), (2, 1, 1), 27, 0.57), ((10, 10, 10), (2, 2, 2), 27, 0.25)])
def test_multigpu_nccl_hierarchy_and_solve(b200, n, grid, stencil, theta):
    """the same comparison with one rank per GPU over NCCL; skipped on boxes with fewer GPUs than ranks"""
    nranks = grid[0] * grid[1] * grid[2]
    if n_gpus() < nranks:
        pytest.skip(f"needs {nranks} GPUs")
    res = run_ranks_nccl(nranks, settings(n, grid, stencil, strong_threshold=theta))
    compare_with_oracle(res, n, grid, stencil, theta)


@pytest.mark.parametrize("method", ["cg", "bicg", "fgmres", "cogmres"])
@pytest.mark.parametrize("transport", ["shm", "nccl"])
def test_multirank_sibling_krylov_methods(b200, method, transport):
    """PCG / BiCGSTAB / FlexGMRES / COGMRES on 2 ranks (all-reduced inner products, distributed preconditioner) against the
    oracle with 2 virtual ranks; the NCCL variant (one GPU per rank, peer-memory halo exchange, graph-replayed coarse
    levels) needs 2 GPUs"""
    n, grid = ((20, 20, 20) if transport == "nccl" else (10, 10, 10)), (2, 1, 1)
    if transport == "nccl" and n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    cfg = settings(n, grid, 7)
    cfg["solver_settings"].update(method=method, kspace=20)
    res = (run_ranks_nccl if transport == "nccl" else run_ranks)(2, cfg)
    o = Oracle(tolerance=1e-8, kspace=20, max_iterations=200)
    o.threads(8)
    o.gen_laplace(n, p=grid, stencil=7)
    o.setup()
    xo = o.solve(method=method)
    x = np.concatenate([d["x"] for d in res])
    its = int(res[0]["iterations"])
    assert all(int(d["iterations"]) == its for d in res)
    assert abs(its - o.iterations) <= 1, (its, o.iterations)
    assert float(res[0]["relres"]) <= max(o.relres * (1 + 1e-6), 1e-8)
    assert np.abs(x - 1.0).max() < 1e-5
    if its == o.iterations:
        assert np.abs(x - xo).max() / np.abs(xo).max() <= 1e-8


@pytest.mark.parametrize("method", ["gmres", "cg"])
def test_multirank_block_jacobi_ilu(b200, method):
    """`preconditioner: ilu` on 2 ranks: every rank factors its own diagonal block (hypre ILU type 0), compared with the
    oracle's 2-virtual-rank block-Jacobi ILU(0)"""
    n, grid = (8, 8, 8), (2, 1, 1)
    cfg = settings(n, grid, 7)
    cfg["solver_settings"].update(method=method, preconditioner="ilu", kspace=40, orthogonalization="mgs")
    res = run_ranks(2, cfg)
    o = Oracle(tolerance=1e-8, kspace=40, max_iterations=200)
    o.gen_laplace(n, p=grid, stencil=7)
    o.ilu_setup()
    xo = o.solve(method=method)
    x = np.concatenate([d["x"] for d in res])
    its = int(res[0]["iterations"])
    assert abs(its - o.iterations) <= 1, (its, o.iterations)
    assert np.abs(x - 1.0).max() < 1e-5
    if its == o.iterations:
        assert np.abs(x - xo).max() / np.abs(xo).max() <= 1e-8


@pytest.mark.parametrize("pmax,tf", [(0, 0.0), (4, 0.15), (0, 0.2)])
def test_multirank_truncation_modes(b200, pmax, tf):
    """the truncation paths of the interpolation kernels with off-rank C points in the rows (P_diag / P_offd re-split after the
    sort): 2 ranks on one GPU against the 2-virtual-rank oracle"""
    n, grid = (10, 10, 10), (2, 1, 1)
    res = run_ranks(2, settings(n, grid, 7, strong_threshold=0.25, p_max_elmts=pmax, trunc_factor=tf))
    compare_with_oracle(res, n, grid, 7, 0.25, p_max_elmts=pmax, trunc_factor=tf)


@pytest.mark.parametrize("n,grid,stencil,theta,pmax", [((10, 10, 10), (2, 1, 1), 7, 0.25, 4), ((6, 6, 6), (2, 2, 1), 27, 0.57, 0),
                                                        ((8, 7, 6), (2, 2, 2), 7, 0.25, 2),
                                                        # found by profiles/parity_sweep_ranks.py: a rank is left without rows on a
                                                        # coarse level; its smoother must still take part in the collective exchanges
                                                        ((8, 5, 5), (2, 2, 2), 27, 0.5, 4)])
def test_multirank_classical_interpolation(b200, n, grid, stencil, theta, pmax):
    """interp_type 0 across ranks: strong F neighbours on other ranks are distributed through their fetched rows (hypre's A_ext),
    C points of other ranks land in P_offd: N ranks on one GPU against the N-virtual-rank oracle"""
    nranks = grid[0] * grid[1] * grid[2]
    res = run_ranks(nranks, settings(n, grid, stencil, strong_threshold=theta, interp_type=0, p_max_elmts=pmax))
    compare_with_oracle(res, n, grid, stencil, theta, interp_type=0, p_max_elmts=pmax)


@pytest.mark.parametrize("n,grid,stencil", [((8, 8, 8), (2, 1, 1), 27), ((8, 8, 8), (2, 2, 1), 27), ((10, 10, 10), (2, 2, 2), 7)])
def test_multirank_reference_halo_map(b200, n, grid, stencil):
    """`halo_map: reference`: the matrix the reference glue really assembles on several ranks (/root/reference/src/HypreSystem.cpp:1553-1556:
    off-rank column = generator-local halo index + (rank == 0 ? iUpper + 1 : 0)) -- nonsymmetric, every rank >= 1 coupled to the first
    rows of rank 0.  Assembled on the device, set up and solved; hierarchy and solution against the oracle with the same surrogate."""
    nranks = grid[0] * grid[1] * grid[2]
    cfg = settings(n, grid, stencil, strong_threshold=0.25)
    cfg["linear_system"]["halo_map"] = "reference"
    res = run_ranks(nranks, cfg)
    compare_with_oracle(res, n, grid, stencil, 0.25, halo_map=1)
    # rank 1's off-rank columns all lie at the start of rank 0's rows
    nloc = n[0] * n[1] * n[2]
    ptr, col = res[1]["A0_ptr"], res[1]["A0_col"]
    off = col[(col < nloc) | (col >= 2 * nloc)]
    assert len(off) > 0 and off.min() == 0 and off.max() == len(np.unique(off)) - 1 and off.max() < nloc


def test_multirank_chebyshev_with_a_rank_without_coarse_rows(b200):
    """found by profiles/parity_sweep_ranks.py: 8 ranks, classical interpolation, Chebyshev smoother, the reference glue's halo
    numbering -- a rank owns no rows of a coarse level; the eigenvalue estimate of the Chebyshev setup must not launch an empty grid
    there and the rank must keep taking part in the inner products"""
    n, grid = (8, 8, 7), (2, 2, 2)
    cfg = settings(n, grid, 27, strong_threshold=0.57, interp_type=0, p_max_elmts=0, relax_type=16)
    cfg["linear_system"]["halo_map"] = "reference"
    res = run_ranks(8, cfg)
    compare_with_oracle(res, n, grid, 27, 0.57, interp_type=0, p_max_elmts=0, relax_type=16, halo_map=1)


@pytest.mark.parametrize("relax", [11, 12])
def test_multirank_two_stage_gauss_seidel(b200, relax):
    """relax 11 / 12 on 2 ranks: the triangle is taken from every rank's own diagonal block (hybrid), off-rank couplings enter
    through the residual; compared with the 2-virtual-rank oracle"""
    n, grid = (12, 12, 12), (2, 1, 1)
    res = run_ranks(2, settings(n, grid, 7, strong_threshold=0.25, relax_type=relax))
    compare_with_oracle(res, n, grid, 7, 0.25, relax_type=relax)


def compare_with_oracle(res, n, grid, stencil, theta, **extra):
    o = Oracle(strong_threshold=theta, **extra)
    o.threads(8)
    o.gen_laplace(n, p=grid, stencil=stencil)
    o.setup()
    xo = o.solve()
    nl = int(res[0]["num_levels"])
    assert all(int(d["num_levels"]) == nl for d in res)
    assert nl == o.num_levels, (nl, o.num_levels)
    for l in range(nl):
        check_csr(stack_csr(res, f"A{l}"), o.matrix(l, "A"), f"A[{l}]")
        if l + 1 < nl:
            cf = np.concatenate([d[f"cf{l}"] for d in res])
            assert np.array_equal(cf, o.cf(l)), f"CF[{l}] differs"
            check_csr(stack_csr(res, f"S{l}"), o.matrix(l, "S"), f"S[{l}]", values=False)
            check_csr(stack_csr(res, f"P{l}"), o.matrix(l, "P"), f"P[{l}]")
            check_csr(stack_csr(res, f"R{l}"), o.matrix(l, "R"), f"R[{l}]")
        np.testing.assert_allclose(np.concatenate([d[f"sm{l}"] for d in res]), o.smoother(l), rtol=1e-13)
    x = np.concatenate([d["x"] for d in res])
    its = int(res[0]["iterations"])
    assert abs(its - o.iterations) <= 1, (its, o.iterations)
    assert float(res[0]["relres"]) <= max(o.relres * 1.5, 1e-8)
    assert np.abs(x - xo).max() <= 1e-6 * max(1.0, np.abs(xo).max())   # both solved to rtol 1e-8
    assert np.abs(x - 1.0).max() < 1e-5                               # generator known answer x == 1


def test_multirank_matches_single_rank_solution(b200):
    """same global problem on 1 and on 4 ranks: the answer is the generator's x == 1 either way"""
    res = run_ranks(4, settings((8, 8, 8), (2, 2, 1), 7, strong_threshold=0.25))
    x = np.concatenate([d["x"] for d in res])
    assert np.abs(x - 1.0).max() < 1e-5
    assert int(res[0]["iterations"]) < 30
