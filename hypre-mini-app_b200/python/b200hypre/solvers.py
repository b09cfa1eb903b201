"""BoomerAMG and ParCSR-GMRES handles over the C ABI (setters mirror /root/reference/src/HypreSystem.cpp:119-326,390-404)."""
import ctypes as C

import numpy as np

from . import F64, I32, I64, VP, _ptr, check, lib


class BoomerAMG:
    # YAML key (boomeramg_settings, HypreSystem.cpp:119-326) -> setter
    _INT = {
        "print_level": "SetPrintLevel", "debug_flag": "SetDebugFlag", "coarsen_type": "SetCoarsenType",
        "cycle_type": "SetCycleType", "relax_type": "SetRelaxType", "num_sweeps": "SetNumSweeps",
        "smooth_num_sweeps": "SetSmoothNumSweeps", "max_iterations": "SetMaxIter", "relax_order": "SetRelaxOrder",
        "max_levels": "SetMaxLevels", "variant": "SetVariant", "rap2": "SetRAP2", "keep_transpose": "SetKeepTranspose",
        "interp_type": "SetInterpType", "min_coarse_size": "SetMinCoarseSize", "max_coarse_size": "SetMaxCoarseSize",
        "pmax_elmts": "SetAggPMaxElmts", "agg_pmax_elmts": "SetAggPMaxElmts", "agg_num_levels": "SetAggNumLevels",
        "agg_interp_type": "SetAggInterpType", "smooth_type": "SetSmoothType", "smooth_num_levels": "SetSmoothNumLevels",
        # not reachable from the mini-app's YAML but part of hypre's API
        "p_max_elmts": "SetPMaxElmts", "cheby_order": "SetChebyOrder", "cheby_eig_est": "SetChebyEigEst",
    }
    _REAL = {
        "tolerance": "SetTol", "strong_threshold": "SetStrongThreshold", "non_galerkin_tol": "SetNonGalerkinTol",
        "trunc_factor": "SetTruncFactor", "max_row_sum": "SetMaxRowSum", "relax_weight": "SetRelaxWt",
        "cheby_fraction": "SetChebyFraction",
    }

    def __init__(self, **settings):
        self.h = VP()
        check(lib().HYPRE_BoomerAMGCreate(C.byref(self.h)), "BoomerAMGCreate")
        self.set(**settings)

    def set(self, **settings):
        L = lib()
        for k, v in settings.items():
            if k in self._INT:
                check(getattr(L, "HYPRE_BoomerAMG" + self._INT[k])(self.h, int(v)), k)
            elif k in self._REAL:
                check(getattr(L, "HYPRE_BoomerAMG" + self._REAL[k])(self.h, float(v)), k)
            elif k in ("down_relax_type", "up_relax_type", "coarse_relax_type"):
                kk = {"down_relax_type": 1, "up_relax_type": 2, "coarse_relax_type": 3}[k]
                check(L.HYPRE_BoomerAMGSetCycleRelaxType(self.h, int(v), kk), k)
            elif k in ("num_down_sweeps", "num_up_sweeps", "num_coarse_sweeps"):
                kk = {"num_down_sweeps": 1, "num_up_sweeps": 2, "num_coarse_sweeps": 3}[k]
                check(L.HYPRE_BoomerAMGSetCycleNumSweeps(self.h, int(v), kk), k)
            else:
                raise KeyError(f"unknown boomeramg setting {k}")

    def setup(self, A, b=None, x=None):
        check(lib().HYPRE_BoomerAMGSetup(self.h, A.object(), b.object() if b else None, x.object() if x else None), "BoomerAMGSetup")

    def solve(self, A, b, x):
        check(lib().HYPRE_BoomerAMGSolve(self.h, A.object(), b.object(), x.object()), "BoomerAMGSolve")

    # ---- inspection -----------------------------------------------------------------------------------------------------
    @property
    def num_levels(self):
        return lib().b200hypre_amg_num_levels(self.h)

    @property
    def setup_seconds(self):
        return lib().b200hypre_amg_setup_seconds(self.h)

    def setup_phases(self):
        names = C.create_string_buffer(1024)
        ms = np.zeros(32)
        n = lib().b200hypre_amg_setup_phases(self.h, names, 1024, _ptr(ms), 32)
        return dict(zip(names.value.decode().split(";"), ms[:n].tolist())) if n else {}

    def level_sizes(self, level, which="A"):
        sizes = np.zeros(4, dtype=np.int64)
        check(lib().b200hypre_amg_level_sizes(self.h, level, ord(which), _ptr(sizes)), "level_sizes")
        return sizes

    def level_matrix(self, level, which="A"):
        """(ptr int32, col int64 global, val float64 or None for 'S')"""
        s = self.level_sizes(level, which)
        ptr = np.zeros(s[0] + 1, dtype=np.int32)
        col = np.zeros(s[2], dtype=np.int64)
        val = None if which == "S" else np.zeros(s[2])
        check(lib().b200hypre_amg_level_get(self.h, level, ord(which), _ptr(ptr), _ptr(col), _ptr(val)), "level_get")
        return ptr, col, val

    def level_cf(self, level):
        n = int(self.level_sizes(level, "A")[0])
        cf = np.zeros(n, dtype=np.int32)
        check(lib().b200hypre_amg_level_cf(self.h, level, _ptr(cf)), "level_cf")
        return cf

    def level_smoother(self, level):
        n = int(self.level_sizes(level, "A")[0])
        d = np.zeros(n)
        check(lib().b200hypre_amg_level_smoother(self.h, level, _ptr(d)), "level_smoother")
        return d

    def level_cheby(self, level):
        out = np.zeros(6)
        check(lib().b200hypre_amg_level_cheby(self.h, level, _ptr(out)))
        return out

    def gs_plan(self, level, which=0, direction=0):
        """wavefront schedule of a Gauss-Seidel pass: dict(built, depth, rows, width, blocks)"""
        out = (C.c_int * 5)()
        check(lib().b200hypre_amg_gs_plan(self.h, int(level), int(which), int(direction), out), "gs_plan")
        return dict(zip(("built", "depth", "rows", "width", "blocks"), [int(v) for v in out]))

    def vcycle_ptr(self, r_ptr, z_ptr):
        check(lib().b200hypre_amg_vcycle(self.h, VP(r_ptr), VP(z_ptr)), "vcycle")

    def apply_as_preconditioner_ptr(self, A, in_ptr, out_ptr):
        """out = M^-1 in through HYPRE_BoomerAMGSolve as a Krylov solver calls it (device pointers)"""
        import ctypes as C
        fn = C.cast(lib().HYPRE_BoomerAMGSolve, VP)
        check(lib().b200hypre_precond_apply(fn, self.h, A.object(), VP(in_ptr), VP(out_ptr)), "precond_apply")

    def relax_ptr(self, level, u_ptr, f_ptr):
        check(lib().b200hypre_amg_relax(self.h, level, VP(u_ptr), VP(f_ptr)), "relax")

    def destroy(self):
        if self.h:
            lib().HYPRE_BoomerAMGDestroy(self.h)
            self.h = None


class ParKrylov:
    """One of hypre's ParCSR Krylov solvers over the C ABI.  `family` is the hypre name part (GMRES, FlexGMRES, COGMRES,
    PCG, BiCGSTAB) selected by the YAML `method` (gmres | fgmres | cogmres | cg | bicg, HypreSystem.cpp:69-85)."""
    FAMILY = {"gmres": "GMRES", "fgmres": "FlexGMRES", "cogmres": "COGMRES", "cg": "PCG", "bicg": "BiCGSTAB"}

    def __init__(self, family="GMRES", tolerance=1e-5, max_iterations=1000, kspace=10, print_level=0, absolute_tolerance=0.0,
                 orthogonalization=None, cgs=None):
        L = lib()
        self.family = family
        self.h = VP()
        for name in ("SetTol", "SetAbsoluteTol"):
            self._fn(name).argtypes = [VP, F64]
        check(self._fn("Create")(0, C.byref(self.h)), family + "Create")
        check(self._fn("SetTol")(self.h, float(tolerance)))
        check(self._fn("SetMaxIter")(self.h, int(max_iterations)))
        if family in ("GMRES", "FlexGMRES", "COGMRES"):  # kspace is commented out for cg / bicg (HypreSystem.cpp:431,448)
            check(self._fn("SetKDim")(self.h, int(kspace)))
        check(self._fn("SetPrintLevel")(self.h, int(print_level)))
        if family == "COGMRES" and cgs is not None:      # HypreSystem.cpp:384
            check(L.HYPRE_ParCSRCOGMRESSetCGS(self.h, int(cgs)))
        if absolute_tolerance:
            check(self._fn("SetAbsoluteTol")(self.h, float(absolute_tolerance)))
        if orthogonalization is not None and family in ("GMRES", "FlexGMRES"):
            check(L.b200hypre_gmres_set_orthogonalization(self.h, {"mgs": 0, "cgs2": 1, "cgs": 2}[orthogonalization]))

    def _fn(self, name):
        return getattr(lib(), f"HYPRE_ParCSR{self.family}{name}")

    def set_precond(self, amg):
        L = lib()
        if amg is None:
            check(self._fn("SetPrecond")(self.h, None, None, None))
        else:  # HypreSystem.cpp:402 with the pointers stored at :323-324 (BoomerAMG) or :367-368 (ILU)
            prefix = getattr(amg, "ABI_PREFIX", "HYPRE_BoomerAMG")
            solve = C.cast(getattr(L, prefix + "Solve"), VP)
            setup = C.cast(getattr(L, prefix + "Setup"), VP)
            check(self._fn("SetPrecond")(self.h, solve, setup, amg.h))

    def setup(self, A, b, x):
        check(self._fn("Setup")(self.h, A.object(), b.object(), x.object()), self.family + "Setup")

    def solve(self, A, b, x):
        rc = self._fn("Solve")(self.h, A.object(), b.object(), x.object())
        if rc == 256:  # HYPRE_ERROR_CONV: reported, not fatal (the mini-app ignores return codes, HypreSystem.cpp:723)
            lib().HYPRE_ClearAllErrors()
            return rc
        check(rc, self.family + "Solve")
        return 0

    @property
    def iterations(self):
        v = I32()
        check(self._fn("GetNumIterations")(self.h, C.byref(v)))
        return v.value

    @property
    def relres(self):
        v = F64()
        check(self._fn("GetFinalRelativeResidualNorm")(self.h, C.byref(v)))
        return v.value

    @property
    def converged(self):
        return lib().b200hypre_solver_converged(self.h)

    @property
    def solve_seconds(self):
        return lib().b200hypre_gmres_solve_seconds(self.h)

    def history(self):
        out = np.zeros(4096)
        n = lib().b200hypre_gmres_get_history(self.h, _ptr(out), 4096)
        return out[:min(n, 4096)].copy()

    def destroy(self):
        if self.h:
            self._fn("Destroy")(self.h)
            self.h = None


class GMRES(ParKrylov):
    def __init__(self, **kw):
        super().__init__("GMRES", **kw)


class AMGSolver:
    """BoomerAMG used as the solver (`method: boomeramg`, HypreSystem.cpp:91-117): same duck type as ParKrylov."""

    def __init__(self, amg, tolerance=1e-5, max_iterations=1000, print_level=0):
        self.amg = amg
        self.tol = tolerance
        amg.set(tolerance=tolerance, max_iterations=max_iterations, print_level=print_level)

    def set_precond(self, amg):
        pass

    def setup(self, A, b, x):
        self.amg.setup(A, b, x)

    def solve(self, A, b, x):
        self.amg.solve(A, b, x)
        return 0

    @property
    def iterations(self):
        v = I32()
        check(lib().HYPRE_BoomerAMGGetNumIterations(self.amg.h, C.byref(v)))
        return v.value

    @property
    def relres(self):
        v = F64()
        check(lib().HYPRE_BoomerAMGGetFinalRelativeResidualNorm(self.amg.h, C.byref(v)))
        return v.value

    @property
    def converged(self):
        return int(self.relres < self.tol)

    def destroy(self):
        pass


class ILU:
    """hypre's ILU handle (block-Jacobi ILU(0) is the built combination): preconditioner (`preconditioner: ilu`, keys of
    ilu_preconditioner_settings, HypreSystem.cpp:328-370) or solver (`method: ilu`, keys of solver_settings, :457-497)."""
    ABI_PREFIX = "HYPRE_ILU"
    _INT = {"ilu_type": "SetType", "max_iterations": "SetMaxIter", "local_reordering": "SetLocalReordering", "print_level": "SetPrintLevel",
            "fill": "SetLevelOfFill", "max_nnz_per_row": "SetMaxNnzPerRow", "iterative_algorithm_type": "SetIterativeSetupType",
            "iterative_setup_option": "SetIterativeSetupOption", "iterative_ilu_max_iterations": "SetIterativeSetupMaxIter",
            "trisolve": "SetTriSolve", "lower_jacobi_iters": "SetLowerJacobiIters", "upper_jacobi_iters": "SetUpperJacobiIters"}
    _REAL = {"tolerance": "SetTol", "drop_threshold": "SetDropThreshold", "iterative_ilu_tolerance": "SetIterativeSetupTolerance"}

    def __init__(self, **settings):
        L = lib()
        for name in self._REAL.values():
            getattr(L, "HYPRE_ILU" + name).argtypes = [VP, F64]
        L.b200hypre_ilu_setup_seconds.restype = F64
        L.b200hypre_ilu_setup_seconds.argtypes = [VP]
        self.h = VP()
        check(L.HYPRE_ILUCreate(C.byref(self.h)), "ILUCreate")
        self.tol = 0.0
        self.set(**settings)

    def set(self, **settings):
        L = lib()
        for k, v in settings.items():
            if k in self._INT:
                check(getattr(L, "HYPRE_ILU" + self._INT[k])(self.h, int(v)), k)
            elif k in self._REAL:
                check(getattr(L, "HYPRE_ILU" + self._REAL[k])(self.h, float(v)), k)
                if k == "tolerance":
                    self.tol = float(v)
            else:
                raise KeyError(f"unknown ilu setting {k}")

    # duck type of ParKrylov so that `method: ilu` plugs into HypreSystem.solve
    def set_precond(self, _):
        pass

    def setup(self, A, b=None, x=None):
        check(lib().HYPRE_ILUSetup(self.h, A.object(), b.object() if b else None, x.object() if x else None), "ILUSetup")

    def solve(self, A, b, x):
        check(lib().HYPRE_ILUSolve(self.h, A.object(), b.object(), x.object()), "ILUSolve")
        return 0

    @property
    def iterations(self):
        v = I32()
        check(lib().HYPRE_ILUGetNumIterations(self.h, C.byref(v)))
        return v.value

    @property
    def relres(self):
        v = F64()
        check(lib().HYPRE_ILUGetFinalRelativeResidualNorm(self.h, C.byref(v)))
        return v.value

    @property
    def converged(self):
        return int(self.tol == 0.0 or self.relres < self.tol)

    @property
    def setup_seconds(self):
        return lib().b200hypre_ilu_setup_seconds(self.h)

    def factors(self, nnz, n):
        val, dinv = np.zeros(nnz), np.zeros(n)
        check(lib().b200hypre_ilu_get_factors(self.h, _ptr(val), _ptr(dinv)), "ilu_get_factors")
        return val, dinv

    def history(self):
        return np.zeros(0)

    num_levels = 0

    def destroy(self):
        if self.h:
            lib().HYPRE_ILUDestroy(self.h)
            self.h = None
