"""ctypes binding of the CPU oracle (oracle/oracle.cpp).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
        if not os.path.exists(path) or os.path.getmtime(path) < max(
                os.path.getmtime(os.path.join(ROOT, "oracle", f)) for f in ("oracle.cpp", "krylov.inc")):
            build()
        L = C.CDLL(path)
        L.ora_create.restype = C.c_void_p
        L.ora_final_relres.restype = C.c_double
        L.ora_setup_seconds.restype = C.c_double
        L.ora_solve_seconds.restype = C.c_double
        for f in ("ora_destroy", "ora_amg_setup", "ora_num_levels", "ora_num_iterations", "ora_converged",
                  "ora_final_relres", "ora_setup_seconds", "ora_solve_seconds"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.ora_set_opt.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
        L.ora_gen_laplace.argtypes = [C.c_void_p] + [C.c_int] * 7
        L.ora_krylov_solve.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double)]
        _LIB = L
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


class Oracle:
    def __init__(self, **opts):
        self.L = lib()
        self.h = C.c_void_p(self.L.ora_create())
        self.set(**opts)

    def __del__(self):
        try:
            self.L.ora_destroy(self.h)
        except Exception:
            pass

    def set(self, **opts):
        for k, v in opts.items():
            rc = self.L.ora_set_opt(self.h, k.encode(), float(v))
            assert rc == 0, f"unknown oracle option {k}"

    def threads(self, n=0):
        return self.L.ora_set_threads(int(n))

    def gen_laplace(self, n, p=(1, 1, 1), stencil=7):
        nx, ny, nz = (n, n, n) if isinstance(n, int) else n
        rc = self.L.ora_gen_laplace(self.h, nx, ny, nz, p[0], p[1], p[2], stencil)
        assert rc == 0
        self.n = nx * ny * nz * p[0] * p[1] * p[2]

    def set_csr(self, ptr, col, val, b, nparts=1, part=None):
        ptr = np.ascontiguousarray(ptr, dtype=np.int32)
        col = np.ascontiguousarray(col, dtype=np.int32)
        val = np.ascontiguousarray(val, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        self.n = len(ptr) - 1
        pp = None if part is None else np.ascontiguousarray(part, dtype=np.int32)
        self.L.ora_set_csr(self.h, C.c_int(self.n), _p(ptr, C.c_int), _p(col, C.c_int), _p(val, C.c_double),
                           _p(b, C.c_double), C.c_int(nparts), _p(pp, C.c_int))

    def rhs(self):
        b = np.empty(self.n)
        self.L.ora_get_rhs(self.h, _p(b, C.c_double))
        return b

    def setup(self):
        self.L.ora_amg_setup(self.h)

    METHODS = {"gmres": 0, "fgmres": 1, "cogmres": 2, "cg": 3, "bicg": 4, "boomeramg": 5, "ilu": 6}

    def solve(self, x0=None, method="gmres"):
        x = np.zeros(self.n) if x0 is None else np.array(x0, dtype=np.float64)
        if method == "gmres":
            self.L.ora_gmres_solve(self.h, _p(x, C.c_double))
        else:
            rc = self.L.ora_krylov_solve(self.h, C.c_int(self.METHODS[method]), _p(x, C.c_double))
            assert rc == 0, f"oracle krylov solve failed ({rc})"
        return x

    def ilu_setup(self):
        """block-Jacobi ILU(0) of the level-0 matrix; becomes the preconditioner of the following solves"""
        assert self.L.ora_ilu_setup(self.h) == 0

    def ilu_apply(self, r):
        r = np.ascontiguousarray(r, dtype=np.float64)
        z = np.zeros_like(r)
        self.L.ora_ilu_apply(self.h, _p(r, C.c_double), _p(z, C.c_double))
        return z

    def ilu_factors(self):
        """(values in the pattern / order of matrix(0, 'A') with u_ii in the diagonal slot, inverted pivots)"""
        nnz = len(self.matrix(0, "A")[1])
        val, dinv = np.zeros(nnz), np.zeros(self.n)
        self.L.ora_ilu_get(self.h, _p(val, C.c_double), _p(dinv, C.c_double))
        return val, dinv

    def vcycle(self, r):
        r = np.ascontiguousarray(r, dtype=np.float64)
        z = np.zeros_like(r)
        self.L.ora_vcycle(self.h, _p(r, C.c_double), _p(z, C.c_double))
        return z

    @property
    def num_levels(self):
        return self.L.ora_num_levels(self.h)

    @property
    def iterations(self):
        return self.L.ora_num_iterations(self.h)

    @property
    def relres(self):
        return self.L.ora_final_relres(self.h)

    @property
    def setup_seconds(self):
        return self.L.ora_setup_seconds(self.h)

    @property
    def solve_seconds(self):
        return self.L.ora_solve_seconds(self.h)

    def history(self):
        out = np.zeros(4096)
        n = self.L.ora_history(self.h, _p(out, C.c_double), C.c_int(4096))
        return out[:n]

    def matrix(self, level, which="A"):
        """returns (ptr, col, val, shape) of level matrix 'A','P','R' or 'S' (val None for S)"""
        sz = np.zeros(3, dtype=np.int64)
        rc = self.L.ora_level_sizes(self.h, C.c_int(level), C.c_int(ord(which)), _p(sz, C.c_longlong))
        assert rc == 0
        nr, nc, nnz = int(sz[0]), int(sz[1]), int(sz[2])
        ptr = np.zeros(nr + 1, dtype=np.int32)
        col = np.zeros(nnz, dtype=np.int32)
        val = np.zeros(nnz) if which != "S" else None
        self.L.ora_level_get(self.h, C.c_int(level), C.c_int(ord(which)), _p(ptr, C.c_int), _p(col, C.c_int), _p(val, C.c_double))
        return ptr, col, val, (nr, nc)

    def scipy(self, level, which="A"):
        import scipy.sparse as sp
        ptr, col, val, shape = self.matrix(level, which)
        if val is None:
            val = np.ones(len(col))
        return sp.csr_matrix((val, col, ptr), shape=shape)

    def cf(self, level):
        nr = self.matrix_rows(level)
        cf = np.zeros(nr, dtype=np.int32)
        self.L.ora_level_cf(self.h, C.c_int(level), _p(cf, C.c_int))
        return cf

    def matrix_rows(self, level):
        sz = np.zeros(3, dtype=np.int64)
        self.L.ora_level_sizes(self.h, C.c_int(level), C.c_int(ord("A")), _p(sz, C.c_longlong))
        return int(sz[0])

    def smoother(self, level):
        out = np.zeros(self.matrix_rows(level))
        self.L.ora_level_smoother(self.h, C.c_int(level), _p(out, C.c_double))
        return out

    def cheby(self, level):
        out = np.zeros(6)
        self.L.ora_level_cheby(self.h, C.c_int(level), _p(out, C.c_double))
        return out

    def measure(self, level):
        out = np.zeros(self.matrix_rows(level))
        rc = self.L.ora_measure(self.h, C.c_int(level), _p(out, C.c_double))
        assert rc == 0
        return out


def rand_stream(seed, n):
    out = np.zeros(n)
    lib().ora_rand_stream(C.c_int(seed), C.c_int(n), _p(out, C.c_double))
    return out
