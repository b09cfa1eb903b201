"""The drop-in boundary checked with the reference's own code (CPU only).

1. The UNMODIFIED reference translation units (/root/reference/src/main.cpp, HypreSystem.cpp), where they lie, pass
   `g++ -fsyntax-only` against include/ (the drop-in headers); with -DB200_NO_MIXED_MINMAX the only diagnostics are the four
   std::min / std::max lines that mix HYPRE_BigInt and HYPRE_Int (the reference's own bug under a mixed-int hypre).
2. oracle/_ref/hypre_app_ref -- those sources linked with libb200hypre.so instead of libHYPRE (oracle/Makefile `ref`) --
   resolves every symbol from the product library, and without a GPU fails loudly instead of computing on the CPU.
3. The fresh MatrixMarket text reader of app/ (mm_text.h) accepts, rejects and sizes files exactly like the reference's
   mmio.c, compiled in place into oracle/_ref/libmmio_ref.so (mmio row).

/root/reference exists only in the build container: tests that need it are skipped elsewhere (the prebuilt oracle/_ref
files travel to the GPU box)."""
import ctypes as C
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
REFDIR = os.path.join(ROOT, "oracle", "_ref")
APP = os.path.join(ROOT, "hypre-mini-app_b200", "app")
LIBDIR = os.path.join(ROOT, "hypre-mini-app_b200", "lib")
have_ref = os.path.isdir(os.path.join(REF, "src"))
needs_ref = pytest.mark.skipif(not have_ref or shutil.which("g++") is None, reason="/root/reference or g++ not present")


def syntax_only(src, *extra):
    cmd = ["g++", "-std=c++14", "-fsyntax-only", "-DUSE_GPU", "-DHYPRE_USING_CUDA", "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(ROOT, "tests", "stubs"), "-I", "/usr/local/cuda/include", *extra, os.path.join(REF, "src", src)]
    p = subprocess.run(cmd, capture_output=True, text=True)
    return [l for l in p.stderr.splitlines() if " error" in l]


@needs_ref
@pytest.mark.parametrize("src", ["main.cpp", "HypreSystem.cpp"])
def test_reference_sources_compile_unmodified_against_include(src):
    errs = syntax_only(src)
    assert errs == [], "\n".join(errs[:20])


@needs_ref
def test_only_required_source_patch_is_the_mixed_int_minmax():
    """without the std::min/max overloads of include/mpi.h the reference needs exactly the patch INTEGRATION.md documents"""
    errs = syntax_only("HypreSystem.cpp", "-DB200_NO_MIXED_MINMAX")
    lines = sorted({int(re.search(r"HypreSystem\.cpp:(\d+):", e).group(1)) for e in errs if "HypreSystem.cpp:" in e})
    assert lines == [1219, 1220, 1291, 1292], errs[:10]
    assert syntax_only("main.cpp", "-DB200_NO_MIXED_MINMAX") == []


@pytest.fixture(scope="module")
def ref_build():
    if have_ref:
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "ref"])
    if not os.path.exists(os.path.join(REFDIR, "hypre_app_ref")):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return REFDIR


def test_reference_host_links_against_product_library(ref_build):
    exe = os.path.join(ref_build, "hypre_app_ref")
    needed = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    assert "libb200hypre.so" in needed and "libHYPRE" not in needed and "libmpi" not in needed
    undefined = subprocess.run(["nm", "-D", "--undefined-only", exe], capture_output=True, text=True).stdout
    wanted = {l.split()[-1].split("@")[0] for l in undefined.splitlines() if re.search(r"\b(HYPRE_|hypre_|MPI_)", l)}
    assert {"HYPRE_BoomerAMGSetup", "HYPRE_ParCSRGMRESSolve", "HYPRE_IJMatrixSetValues2", "MPI_Comm_split_type"} <= wanted
    exported = subprocess.run(["nm", "-D", "--defined-only", os.path.join(LIBDIR, "libb200hypre.so")], capture_output=True, text=True).stdout
    have = {l.split()[-1] for l in exported.splitlines()}
    assert wanted <= have, f"reference needs symbols the library does not export: {sorted(wanted - have)}"


def test_reference_host_fails_loudly_without_gpu(ref_build, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by tests/test_gpu_app.py::test_unmodified_reference_host_runs")
    y = tmp_path / "in.yaml"
    y.write_text("linear_system:\n  type: build_27pt_stencil\n  nx: 4\n  ny: 4\n  nz: 4\nsolver_settings:\n  method: gmres\n"
                 "  preconditioner: boomeramg\nboomeramg_settings:\n  coarsen_type: 8\n")
    p = subprocess.run([os.path.join(ref_build, "hypre_app_ref"), str(y)], capture_output=True, text=True, timeout=120)
    assert "no CPU fallback" in (p.stdout + p.stderr)


# ---- mmio cross-check -------------------------------------------------------------------------------------------------------
BANNERS = [
    "%%MatrixMarket matrix coordinate real general",
    "%%MatrixMarket MATRIX Coordinate REAL General",        # fields are case-insensitive
    "%%matrixmarket matrix coordinate real general",        # the banner token is not
    "%%MatrixMarket matrix coordinate real symmetric",
    "%%MatrixMarket matrix coordinate real skew-symmetric",
    "%%MatrixMarket matrix coordinate real hermitian",      # invalid combination (mm_is_valid)
    "%%MatrixMarket matrix coordinate complex hermitian",
    "%%MatrixMarket matrix coordinate complex general",
    "%%MatrixMarket matrix coordinate pattern general",
    "%%MatrixMarket matrix coordinate pattern symmetric",
    "%%MatrixMarket matrix coordinate pattern skew-symmetric",  # invalid
    "%%MatrixMarket matrix coordinate integer general",
    "%%MatrixMarket matrix coordinate double general",      # not a MatrixMarket field
    "%%MatrixMarket matrix array real general",
    "%%MatrixMarket matrix array pattern general",          # invalid
    "%%MatrixMarket vector coordinate real general",
    "%%MatrixMarket matrix coordinate real",                # too few tokens
    "%%MatrixMarketX matrix coordinate real general",       # prefix match
    "%MatrixMarket matrix coordinate real general",
    "  %%MatrixMarket   matrix\tcoordinate  real   general  ",
]
BODIES = [
    "3 3 2\n1 1 1.0\n2 2 2.0\n",
    "% comment\n%another\n4 5 1\n1 1 1.0\n",
    "\n\n7 8 9\n1 1 1.0\n",                                  # blank lines before the sizes
    "%c\n\n   10   11   12   \n",
]


@pytest.fixture(scope="module")
def mm_tools(ref_build, tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("bin") / "mm_probe")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", APP, os.path.join(ROOT, "tests", "cpp", "mm_probe.cpp"), "-o", exe])
    lib = C.CDLL(os.path.join(ref_build, "libmmio_ref.so"))
    libc = C.CDLL(None)
    libc.fopen.restype = C.c_void_p
    libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
    libc.fclose.argtypes = [C.c_void_p]
    lib.mm_read_banner.argtypes = [C.c_void_p, C.c_char_p]
    lib.mm_read_mtx_crd_size.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.mm_read_mtx_array_size.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.mm_is_valid.argtypes = [C.c_char_p]
    return exe, lib, libc


def reference_reading(lib, libc, path):
    """what /root/reference/src/HypreSystem.cpp:1683-1690 does with the file"""
    fh = libc.fopen(path.encode(), b"r")
    code = C.create_string_buffer(4)
    try:
        if lib.mm_read_banner(fh, code) != 0:
            return "error"
        tc = code.raw.decode()
        if not lib.mm_is_valid(code):
            return "error"
        m, n, nz = C.c_int(), C.c_int(), C.c_int()
        if tc[1] == "C":
            if lib.mm_read_mtx_crd_size(fh, C.byref(m), C.byref(n), C.byref(nz)) != 0:
                return "error"
        else:
            if lib.mm_read_mtx_array_size(fh, C.byref(m), C.byref(n)) != 0:
                return "error"
        return f"ok {tc[1]}{tc[2]}{tc[3]} {m.value} {n.value} {nz.value}"
    finally:
        libc.fclose(fh)


@pytest.mark.parametrize("eol", ["\n", "\r\n"])
def test_mm_text_reader_agrees_with_reference_mmio(mm_tools, tmp_path, eol):
    exe, lib, libc = mm_tools
    checked = 0
    for bi, banner in enumerate(BANNERS):
        for ci, body in enumerate(BODIES):
            if "array" in banner:
                body = body.replace("3 3 2", "3 3").replace("4 5 1", "4 5").replace("7 8 9", "7 8").replace("10   11   12", "10   11")
            f = tmp_path / f"m{bi}_{ci}.mtx"
            f.write_bytes((banner + "\n" + body).replace("\n", eol).encode())
            ref = reference_reading(lib, libc, str(f))
            got = subprocess.run([exe, str(f)], capture_output=True, text=True).stdout.strip()
            if ref == "error":
                assert got.startswith("error"), f"{banner!r}: reference rejects, product says {got!r}"
            else:
                assert got == ref, f"{banner!r} / body {ci}: reference {ref!r}, product {got!r}"
            checked += 1
    assert checked == len(BANNERS) * len(BODIES)
