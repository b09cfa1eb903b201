"""ctypes binding of libb200hypre.so (the C ABI declared in include/HYPRE.h and include/b200hypre.h).

Harness only: tests/, bench.py and __graft_entry__.py drive the product through this module exactly the way
/root/reference/src/HypreSystem.cpp drives hypre (IJ create/set/assemble -> SetPrecond -> Setup -> Solve,
HypreSystem.cpp:546-636,673-737).  There is no Python compute path and no CPU fallback: every call lands in the CUDA
library, which fails loudly when no sm_100a device is present.
"""
import ctypes as C
import os
import subprocess

import numpy as np

PKG_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB_PATH = os.path.join(PKG_ROOT, "lib", "libb200hypre.so")
HYPRE_PARCSR = 5555
HYPRE_MEMORY_HOST, HYPRE_MEMORY_DEVICE = 0, 1

_lib = None

I32, I64, F64, VP = C.c_int, C.c_longlong, C.c_double, C.c_void_p
PI32, PI64, PF64 = C.POINTER(C.c_int), C.POINTER(C.c_longlong), C.POINTER(C.c_double)


class HypreError(RuntimeError):
    pass


def build(force=False):
    """compile libb200hypre.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)"""
    if force:
        subprocess.check_call(["make", "-s", "-C", PKG_ROOT, "clean"])
    subprocess.check_call(["make", "-s", "-j8", "-C", PKG_ROOT], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise HypreError(f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no fallback path)")
        L = C.CDLL(LIB_PATH)
        L.b200hypre_last_error.restype = C.c_char_p
        L.b200hypre_stream.restype = VP
        L.b200hypre_vector_data.restype = VP
        L.b200hypre_vector_data.argtypes = [VP]
        L.b200hypre_vector_local_size.argtypes = [VP]
        L.b200hypre_launch_count.restype = C.c_ulonglong
        L.b200hypre_amg_setup_seconds.restype = F64
        L.b200hypre_amg_setup_seconds.argtypes = [VP]
        L.b200hypre_gmres_solve_seconds.restype = F64
        L.b200hypre_gmres_solve_seconds.argtypes = [VP]
        L.b200hypre_spmv_kernel_name.restype = C.c_char_p
        L.b200hypre_spmv_kernel_name.argtypes = [VP]
        L.b200hypre_malloc.restype = VP
        L.b200hypre_malloc.argtypes = [C.c_size_t, I32]
        L.b200hypre_free.argtypes = [VP, I32]
        L.b200hypre_memcpy.argtypes = [VP, VP, C.c_size_t, I32, I32]
        L.b200hypre_host_free.argtypes = [VP]
        L.HYPRE_IJMatrixCreate.argtypes = [I32, I64, I64, I64, I64, C.POINTER(VP)]
        L.HYPRE_IJMatrixSetValues2.argtypes = [VP, I32, VP, VP, VP, VP, VP]
        L.HYPRE_IJMatrixAddToValues2.argtypes = [VP, I32, VP, VP, VP, VP, VP]
        L.HYPRE_IJVectorCreate.argtypes = [I32, I64, I64, C.POINTER(VP)]
        L.HYPRE_IJVectorSetValues.argtypes = [VP, I32, VP, VP]
        L.HYPRE_IJVectorGetValues.argtypes = [VP, I32, VP, VP]
        L.HYPRE_ParCSRGMRESSetTol.argtypes = [VP, F64]
        L.HYPRE_ParCSRGMRESSetAbsoluteTol.argtypes = [VP, F64]
        L.HYPRE_ParCSRMatrixMatvec.argtypes = [F64, VP, VP, F64, VP]
        L.HYPRE_ParVectorSetConstantValues.argtypes = [VP, F64]
        for name in ("SetTol", "SetStrongThreshold", "SetMaxRowSum", "SetTruncFactor", "SetRelaxWt", "SetChebyFraction",
                     "SetNonGalerkinTol"):
            getattr(L, "HYPRE_BoomerAMG" + name).argtypes = [VP, F64]
        L.b200hypre_amg_relax.argtypes = [VP, I32, VP, VP]
        L.b200hypre_amg_vcycle.argtypes = [VP, VP, VP]
        _lib = L
    return _lib


def check(rc, what=""):
    if rc != 0:
        L = lib()
        msg = L.b200hypre_last_error().decode()
        L.HYPRE_ClearAllErrors()
        raise HypreError(f"{what} failed (flag {rc}): {msg}")


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


# ---- ranks ---------------------------------------------------------------------------------------------------------------
def comm_init(rank=0, nranks=1, nccl_id=None):
    L = lib()
    buf = None if nccl_id is None else (C.c_char * 128).from_buffer_copy(bytes(nccl_id))
    check(L.b200hypre_comm_init(rank, nranks, buf), "comm_init")


def comm_init_shm(rank, nranks, name, slot_bytes=0):
    """N ranks sharing one GPU (NCCL refuses that): host-staged shared-memory transport, used by the N-rank parity tests"""
    L = lib()
    L.b200hypre_comm_init_shm.argtypes = [I32, I32, C.c_char_p, C.c_size_t]
    check(L.b200hypre_comm_init_shm(rank, nranks, name.encode(), slot_bytes), "comm_init_shm")


def get_unique_id():
    buf = (C.c_char * 128)()
    check(lib().b200hypre_get_unique_id(buf), "get_unique_id")
    return bytes(buf)


def comm_init_torch():
    """one process per GPU under torchrun: ship the NCCL id through torch.distributed, then wire the library"""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    if world == 1:
        comm_init(0, 1)
        return
    if rank == 0:
        ident = torch.tensor(list(get_unique_id()), dtype=torch.uint8)
    else:
        ident = torch.zeros(128, dtype=torch.uint8)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    ident = ident.to(dev)
    dist.broadcast(ident, 0)
    comm_init(rank, world, bytes(ident.cpu().tolist()))


def process_grid(nprocs):
    px, py, pz = I32(), I32(), I32()
    lib().b200hypre_process_grid(nprocs, C.byref(px), C.byref(py), C.byref(pz))
    return px.value, py.value, pz.value


def block_partition(n, nranks):
    out = np.zeros(nranks + 1, dtype=np.int64)
    lib().b200hypre_block_partition(I64(n), nranks, _ptr(out))
    return out


def plan_recv(col_map, col_starts):
    col_map = np.ascontiguousarray(col_map, dtype=np.int64)
    col_starts = np.ascontiguousarray(col_starts, dtype=np.int64)
    nr = len(col_starts) - 1
    rr = np.zeros(nr, dtype=np.int32)
    ro = np.zeros(nr + 1, dtype=np.int32)
    k = lib().b200hypre_plan_recv(_ptr(col_map), len(col_map), _ptr(col_starts), nr, _ptr(rr), _ptr(ro))
    return rr[:k].copy(), ro[:k + 1].copy()


def halo_transport():
    import ctypes as C
    f = lib().b200hypre_halo_transport
    f.restype = C.c_char_p
    return f().decode()


def launch_count():
    return int(lib().b200hypre_launch_count())


def pool_stats():
    """bytes of the library's device memory pool: reserved now / high-water, in use now / high-water"""
    out = (C.c_ulonglong * 4)()
    check(lib().b200hypre_pool_stats(out), "pool_stats")
    return {"reserved": int(out[0]), "reserved_high": int(out[1]), "used": int(out[2]), "used_high": int(out[3])}


def pool_reserve(nbytes):
    L = lib()
    L.b200hypre_pool_reserve.argtypes = [C.c_ulonglong]
    check(L.b200hypre_pool_reserve(int(nbytes)), "pool_reserve")


def profile_enable(on=True, min_bytes=0.0):
    L = lib()
    L.b200hypre_profile_enable.argtypes = [I32, F64]
    check(L.b200hypre_profile_enable(1 if on else 0, float(min_bytes)), "profile_enable")


def profile_summary():
    """{class: (total_ms, calls, algorithmic_bytes)} of the scopes recorded since profile_enable"""
    names = C.create_string_buffer(4096)
    ms, calls, nbytes = np.zeros(64), np.zeros(64), np.zeros(64)
    n = lib().b200hypre_profile_summary(names, 4096, _ptr(ms), _ptr(calls), _ptr(nbytes), 64)
    keys = names.value.decode().split(";") if n else []
    return {k: (float(ms[i]), int(calls[i]), float(nbytes[i])) for i, k in enumerate(keys)}


def stream_ptr():
    return lib().b200hypre_stream()


def synchronize():
    check(lib().b200hypre_device_synchronize(), "synchronize")


# ---- handles -------------------------------------------------------------------------------------------------------------
class IJMatrix:
    def __init__(self, ilower, iupper, jlower=None, jupper=None):
        L = lib()
        self.h = VP()
        jlower = ilower if jlower is None else jlower
        jupper = iupper if jupper is None else jupper
        check(L.HYPRE_IJMatrixCreate(0, ilower, iupper, jlower, jupper, C.byref(self.h)), "IJMatrixCreate")
        check(L.HYPRE_IJMatrixSetObjectType(self.h, HYPRE_PARCSR))
        check(L.HYPRE_IJMatrixInitialize(self.h), "IJMatrixInitialize")
        self.ilower, self.iupper = ilower, iupper

    def set_values_ptr(self, n, rows, cols, vals, add=False):
        """COO triplets, one entry per 'row' (ncols == NULL), raw host or device addresses (HypreSystem.cpp:942)"""
        f = lib().HYPRE_IJMatrixAddToValues2 if add else lib().HYPRE_IJMatrixSetValues2
        check(f(self.h, int(n), None, VP(rows), None, VP(cols), VP(vals)), "IJMatrixSetValues2")

    def set_values(self, rows, cols, vals, add=False):
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        cols = np.ascontiguousarray(cols, dtype=np.int64)
        vals = np.ascontiguousarray(vals, dtype=np.float64)
        self.set_values_ptr(len(rows), rows.ctypes.data, cols.ctypes.data, vals.ctypes.data, add)

    def assemble(self):
        check(lib().HYPRE_IJMatrixAssemble(self.h), "IJMatrixAssemble")

    def object(self):
        o = VP()
        check(lib().HYPRE_IJMatrixGetObject(self.h, C.byref(o)))
        return o

    def print(self, filename):
        check(lib().HYPRE_IJMatrixPrint(self.h, filename.encode()), "IJMatrixPrint")

    def to_csr(self):
        return parcsr_to_host(self.object())

    def destroy(self):
        if self.h:
            lib().HYPRE_IJMatrixDestroy(self.h)
            self.h = None


class IJVector:
    def __init__(self, jlower, jupper):
        L = lib()
        self.h = VP()
        check(L.HYPRE_IJVectorCreate(0, jlower, jupper, C.byref(self.h)), "IJVectorCreate")
        check(L.HYPRE_IJVectorSetObjectType(self.h, HYPRE_PARCSR))
        check(L.HYPRE_IJVectorInitialize(self.h), "IJVectorInitialize")
        self.jlower, self.jupper = jlower, jupper
        self.n = jupper - jlower + 1

    def set_values_ptr(self, n, idx, vals):
        check(lib().HYPRE_IJVectorSetValues(self.h, int(n), VP(idx) if idx else None, VP(vals)), "IJVectorSetValues")

    def set_values(self, vals, idx=None):
        vals = np.ascontiguousarray(vals, dtype=np.float64)
        if idx is None:
            idx = np.arange(self.jlower, self.jlower + len(vals), dtype=np.int64)
        idx = np.ascontiguousarray(idx, dtype=np.int64)
        self.set_values_ptr(len(vals), idx.ctypes.data, vals.ctypes.data)

    def assemble(self):
        check(lib().HYPRE_IJVectorAssemble(self.h))

    def get_values(self):
        out = np.zeros(self.n)
        check(lib().HYPRE_IJVectorGetValues(self.h, self.n, None, _ptr(out)), "IJVectorGetValues")
        return out

    def get_values_ptr(self, out_ptr):
        """all local values into a caller buffer (host, e.g. pinned, or device address)"""
        check(lib().HYPRE_IJVectorGetValues(self.h, self.n, None, VP(out_ptr)), "IJVectorGetValues")

    def object(self):
        o = VP()
        check(lib().HYPRE_IJVectorGetObject(self.h, C.byref(o)))
        return o

    def fill(self, value):
        check(lib().HYPRE_ParVectorSetConstantValues(self.object(), float(value)))

    def data_ptr(self):
        return lib().b200hypre_vector_data(self.object())

    def print(self, filename):
        check(lib().HYPRE_IJVectorPrint(self.h, filename.encode()), "IJVectorPrint")

    def destroy(self):
        if self.h:
            lib().HYPRE_IJVectorDestroy(self.h)
            self.h = None


def parcsr_to_host(par):
    """(ptr int32, col int64 global, val float64, sizes) of the local rows"""
    L = lib()
    sizes = np.zeros(4, dtype=np.int64)
    check(L.b200hypre_parcsr_sizes(par, _ptr(sizes)))
    ptr = np.zeros(sizes[0] + 1, dtype=np.int32)
    col = np.zeros(sizes[2], dtype=np.int64)
    val = np.zeros(sizes[2])
    check(L.b200hypre_parcsr_get(par, _ptr(ptr), _ptr(col), _ptr(val)))
    return ptr, col, val, sizes


def matvec(A, x, y, alpha=1.0, beta=0.0):
    check(lib().HYPRE_ParCSRMatrixMatvec(alpha, A.object(), x.object(), beta, y.object()), "Matvec")


from .solvers import ILU, AMGSolver, BoomerAMG, GMRES, ParKrylov  # noqa: E402
from .system import HypreSystem, generate_laplace  # noqa: E402
