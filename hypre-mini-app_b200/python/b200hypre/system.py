"""HypreSystem: harness-side mirror of nalu::HypreSystem (/root/reference/src/HypreSystem.h:66-298) over the C ABI.

Same life cycle and the same timer names as the reference: setup_precon_and_solver() -> load() -> solve()
(/root/reference/src/main.cpp:172-178), "Preconditioner setup" = SetPrecond + GMRESSetup (which runs the whole
BoomerAMG setup), "Solve" = GMRESSolve (HypreSystem.cpp:685-696,723-734).  The settings dict has the layout of
etc/hypre_app.yaml: {"linear_system": {...}, "solver_settings": {...}, "boomeramg_settings": {...}}.
The stand-alone C++ twin lives in hypre-mini-app_b200/app/.
"""
import ctypes as C
import time

import numpy as np

from . import (HYPRE_MEMORY_DEVICE, I32, I64, VP, HypreError, IJMatrix, IJVector, check, lib, process_grid, synchronize)
from .solvers import ILU, AMGSolver, BoomerAMG, ParKrylov


class DeviceCOO:
    """device arrays produced by the generator; freed explicitly (caller-owned, HypreSystem.cpp:626-633)"""

    def __init__(self, rows, cols, vals, rhs, nnz, n_local, row_start, global_rows):
        self.rows, self.cols, self.vals, self.rhs = rows, cols, vals, rhs
        self.nnz, self.n_local, self.row_start, self.global_rows = nnz, n_local, row_start, global_rows

    def free(self):
        L = lib()
        for p in (self.rows, self.cols, self.vals, self.rhs):
            if p:
                L.b200hypre_free(p, HYPRE_MEMORY_DEVICE)
        self.rows = self.cols = self.vals = self.rhs = None


def generate_laplace(nx, ny, nz, grid=(1, 1, 1), rank=0, stencil=7):
    """laplace_3d_weak_scaling generator on the device (laplace_3d_weak_scaling.hpp:604-1090 semantics)"""
    L = lib()
    rows, cols, vals, rhs = VP(), VP(), VP(), VP()
    nnz, nloc, rs, gr = I64(), I64(), I64(), I64()
    check(L.b200hypre_generate_laplace(nx, ny, nz, grid[0], grid[1], grid[2], rank, stencil, C.byref(rows), C.byref(cols),
                                       C.byref(vals), C.byref(rhs), C.byref(nnz), C.byref(nloc), C.byref(rs), C.byref(gr)),
          "generate_laplace")
    return DeviceCOO(rows.value, cols.value, vals.value, rhs.value, nnz.value, nloc.value, rs.value, gr.value)


def apply_reference_halo_map(coo, rank):
    """`halo_map: reference`: off-rank columns of the device COO become generator-local halo index + (rank == 0 ? iUpper + 1 : 0),
    what /root/reference/src/HypreSystem.cpp:1553-1556 hands to hypre; returns the number of halo points"""
    L = lib()
    L.b200hypre_apply_reference_halo_map.argtypes = [VP, I64, I64, I64, I32, C.POINTER(I64)]
    nh = I64()
    check(L.b200hypre_apply_reference_halo_map(coo.cols, coo.nnz, coo.row_start, coo.row_start + coo.n_local - 1, rank, C.byref(nh)),
          "apply_reference_halo_map")
    return nh.value


WORKLOAD_KINDS = {"poisson27_hex": (0, 20260, 1), "momentum3": (1, 20261, 3)}   # kind id, seed, rows per node


def generate_workload(kind, n, i_lower, i_upper):
    """rows [i_lower, i_upper] of the synthetic systems of BASELINE.json configs[3] / [4], built on the device
    (csrc/workload_rows.cuh; tests/workloads.py holds the numpy builders of the same matrices)"""
    L = lib()
    kid, seed, _ = WORKLOAD_KINDS[kind]
    rows, cols, vals, rhs = VP(), VP(), VP(), VP()
    nnz, gr = I64(), I64()
    L.b200hypre_generate_workload.argtypes = [I32, I32, I64, I64, C.c_ulonglong, C.POINTER(VP), C.POINTER(VP), C.POINTER(VP), C.POINTER(VP),
                                              C.POINTER(I64), C.POINTER(I64)]
    check(L.b200hypre_generate_workload(kid, n, i_lower, i_upper, seed, C.byref(rows), C.byref(cols), C.byref(vals), C.byref(rhs),
                                        C.byref(nnz), C.byref(gr)), "generate_workload")
    return DeviceCOO(rows.value, cols.value, vals.value, rhs.value, nnz.value, i_upper - i_lower + 1, i_lower, gr.value)


DEFAULT_AMG = dict(print_level=0, debug_flag=0, coarsen_type=8, cycle_type=1, relax_type=18, num_sweeps=1, smooth_num_sweeps=1,
                   tolerance=0.0, max_iterations=1, relax_order=0, max_levels=20, strong_threshold=0.25, interp_type=6)


class HypreSystem:
    def __init__(self, settings, rank=0, nranks=1):
        self.cfg = settings
        self.rank, self.nranks = rank, nranks
        self.timers = []
        self.A = self.b = self.x = None
        self.solver = self.precond = self.amg_solver = None
        self.n_local = self.i_lower = self.i_upper = self.global_rows = 0

    def _timer(self, name, t0):
        self.timers.append((name, time.perf_counter() - t0))

    # -- HypreSystem.cpp:49-89 ------------------------------------------------------------------------------------------
    def setup_precon_and_solver(self):
        ss = self.cfg.get("solver_settings", {})
        method = ss.get("method", "gmres")
        if method not in ParKrylov.FAMILY and method not in ("boomeramg", "ilu"):
            raise ValueError(f"Invalid option for solver method provided: {method}")
        precond = ss.get("preconditioner", "boomeramg")
        if precond == "boomeramg":
            amg = dict(DEFAULT_AMG)
            amg.update(self.cfg.get("boomeramg_settings", {}))
            self.precond = BoomerAMG(**amg)
        elif precond == "ilu":   # setup_ilu_precond, HypreSystem.cpp:328-370 (defaults of the reference)
            ilu = dict(ilu_type=0, max_iterations=1, tolerance=0.0, local_reordering=0, print_level=0, fill=0, trisolve=1)
            ilu.update(self.cfg.get("ilu_preconditioner_settings", {}))
            self.precond = ILU(**ilu)
        elif precond == "none":
            self.precond = None
        else:
            raise ValueError(f"Invalid option for preconditioner provided {precond}")
        if method == "ilu":       # setup_ilu, HypreSystem.cpp:457-497: ILU sweeps as the solver, no preconditioner
            keys = set(ILU._INT) | set(ILU._REAL)
            ilu = dict(ilu_type=0, max_iterations=20, tolerance=0.0, local_reordering=0, print_level=0, fill=0, trisolve=1)
            ilu.update({k: v for k, v in ss.items() if k in keys})
            if self.precond is not None:
                self.precond.destroy()
            self.precond = None
            self.solver = ILU(**ilu)
            return
        if method == "boomeramg":
            # setup_boomeramg_solver (HypreSystem.cpp:91-117): BoomerAMG is the solver, no preconditioner.  The reference
            # applies the coarsening / relaxation keys of solver_settings to precond_ (:100-110), leaving solver_ on hypre's
            # defaults; here they configure the solver itself, which is what those lines evidently intend.
            amg = dict(DEFAULT_AMG)
            amg.update(coarsen_type=ss.get("coarsen_type", 8), cycle_type=ss.get("cycle_type", 1), relax_type=ss.get("relax_type", 6),
                       num_sweeps=ss.get("num_sweeps", 1), smooth_num_sweeps=ss.get("smooth_num_sweeps", 1),
                       relax_order=ss.get("relax_order", 1), max_levels=ss.get("max_levels", 20),
                       strong_threshold=ss.get("strong_threshold", 0.57))
            for k in ("interp_type", "p_max_elmts", "trunc_factor"):
                if k in ss:
                    amg[k] = ss[k]
            if self.precond is not None:
                self.precond.destroy()
            self.precond = None
            self.amg_solver = BoomerAMG(**amg)
            self.solver = AMGSolver(self.amg_solver, tolerance=ss.get("tolerance", 1e-5), max_iterations=ss.get("max_iterations", 1000),
                                    print_level=ss.get("print_level", 0))
            return
        self.solver = ParKrylov(ParKrylov.FAMILY[method], tolerance=ss.get("tolerance", 1e-5), max_iterations=ss.get("max_iterations", 1000),
                                kspace=ss.get("kspace", 10), print_level=ss.get("print_level", 0),
                                orthogonalization=ss.get("orthogonalization"), cgs=ss.get("cgs", 0))

    # -- HypreSystem.cpp:16-47 --------------------------------------------------------------------------------------------
    def load(self):
        ls = self.cfg.get("linear_system", {})
        kind = ls.get("type", "matrix_market")
        if kind in ("build_27pt_stencil", "build_7pt_stencil"):
            self._load_generator(ls, 27 if kind == "build_27pt_stencil" else 7)
        elif kind in WORKLOAD_KINDS:
            self._load_workload(ls, kind)
        elif kind == "csr":
            self._load_csr(ls)
        elif kind == "coo":
            self._load_coo(ls)
        else:
            raise ValueError(f"linear_system type '{kind}': use the C++ hypre_app for file formats")

    def _init_system(self, i_lower, i_upper):
        t0 = time.perf_counter()
        self.i_lower, self.i_upper = i_lower, i_upper
        self.n_local = i_upper - i_lower + 1
        self.A = IJMatrix(i_lower, i_upper)
        self.b = IJVector(i_lower, i_upper)
        self.x = IJVector(i_lower, i_upper)
        self.x.fill(0.0)
        self._timer("Initialize system", t0)

    def _load_generator(self, ls, stencil):
        nx, ny, nz = ls.get("nx", 128), ls.get("ny", 128), ls.get("nz", 128)
        grid = tuple(ls.get("grid", process_grid(self.nranks)))
        t0 = time.perf_counter()
        coo = generate_laplace(nx, ny, nz, grid, self.rank, stencil)
        halo_map = ls.get("halo_map", "true")
        if halo_map == "reference":      # the reference glue's surrogate off-rank numbering (HypreSystem.cpp:1553-1556)
            self.n_halo = apply_reference_halo_map(coo, self.rank)
        elif halo_map != "true":
            raise HypreError(f"linear_system.halo_map must be 'true' or 'reference', got {halo_map!r}")
        self.global_rows = coo.global_rows
        self._init_system(coo.row_start, coo.row_start + coo.n_local - 1)
        self.A.set_values_ptr(coo.nnz, coo.rows, coo.cols, coo.vals)
        self.b.set_values_ptr(coo.n_local, 0, coo.rhs)
        synchronize()
        coo.free()
        self._timer("Build 27Pt Stencil HYPRE matrix", t0)

    def _load_workload(self, ls, kind):
        """device-generated poisson27_hex / momentum3 system of `n`^3 nodes, block rows per rank (HypreSystem.cpp:525-544)"""
        from . import block_partition
        n = int(ls["n"])
        g_rows = WORKLOAD_KINDS[kind][2] * n ** 3
        part = block_partition(g_rows, self.nranks)
        lo, hi = int(part[self.rank]), int(part[self.rank + 1]) - 1
        t0 = time.perf_counter()
        coo = generate_workload(kind, n, lo, hi)
        self.global_rows = g_rows
        self._init_system(lo, hi)
        self.A.set_values_ptr(coo.nnz, coo.rows, coo.cols, coo.vals)
        self.b.set_values_ptr(coo.n_local, 0, coo.rhs)
        synchronize()
        coo.free()
        self._timer("Build HYPRE matrix", t0)

    def _load_coo(self, ls):
        """COO triplets (global ids, one entry per 'row') + rhs given as raw HOST or DEVICE addresses: exactly what
        hypre_matrix_set_values / hypre_vector_set_values hand to the IJ interface (HypreSystem.cpp:897-1015)"""
        self.global_rows = ls["global_rows"]
        self._init_system(ls["row_start"], ls["row_start"] + ls["n_local"] - 1)
        t0 = time.perf_counter()
        self.A.set_values_ptr(ls["nnz"], ls["rows"], ls["cols"], ls["vals"])
        self._timer("Build HYPRE matrix", t0)
        t0 = time.perf_counter()
        self.b.set_values_ptr(ls["n_local"], 0, ls["rhs"])
        self._timer("Build HYPRE vector", t0)

    def _load_csr(self, ls):
        """global CSR given as host arrays (tests): every rank keeps its block rows (HypreSystem.cpp:525-544)"""
        from . import block_partition
        ptr, col, val, rhs = ls["ptr"], ls["col"], ls["val"], ls["rhs"]
        n = len(ptr) - 1
        self.global_rows = n
        part = ls.get("partition")
        part = block_partition(n, self.nranks) if part is None else np.asarray(part, dtype=np.int64)
        lo, hi = int(part[self.rank]), int(part[self.rank + 1]) - 1
        self._init_system(lo, hi)
        t0 = time.perf_counter()
        b, e = ptr[lo], ptr[hi + 1]
        rows = np.repeat(np.arange(lo, hi + 1, dtype=np.int64), np.diff(ptr[lo:hi + 2]))
        self.A.set_values(rows, np.asarray(col[b:e], dtype=np.int64), val[b:e])
        self._timer("Build HYPRE matrix", t0)
        t0 = time.perf_counter()
        self.b.set_values(np.asarray(rhs[lo:hi + 1], dtype=np.float64))
        self._timer("Build HYPRE vector", t0)

    # -- HypreSystem.cpp:600-636 ------------------------------------------------------------------------------------------
    def assemble_system(self):
        t0 = time.perf_counter()
        self.A.assemble()
        self.b.assemble()
        self.x.assemble()
        synchronize()
        self._timer("Assemble system", t0)

    # -- HypreSystem.cpp:673-737 ------------------------------------------------------------------------------------------
    def solve(self):
        from . import lib
        self.assemble_system()
        L = lib()
        if self.nranks > 1:
            L.b200hypre_barrier()
        t0 = time.perf_counter()
        self.solver.set_precond(self.precond)
        self.solver.setup(self.A, self.b, self.x)
        synchronize()
        if self.nranks > 1:
            L.b200hypre_barrier()
        self._timer("Preconditioner setup", t0)
        t0 = time.perf_counter()
        self.solver.solve(self.A, self.b, self.x)
        synchronize()
        if self.nranks > 1:
            L.b200hypre_barrier()
        self._timer("Solve", t0)

    def timer(self, name):
        return sum(t for n, t in self.timers if n == name)

    def solution(self):
        return self.x.get_values()

    def destroy_system(self):
        for o in (self.solver, self.precond, getattr(self, "amg_solver", None), self.A, self.b, self.x):
            if o is not None:
                o.destroy()
        self.solver = self.precond = self.amg_solver = self.A = self.b = self.x = None
