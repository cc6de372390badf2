"""CPU: the C-ABI shared object loads without a GPU and exports every symbol the header declares;
host-side argument handling of the python mirror.  No compute calls."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "accblas_b200.h")


def header_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(accblas_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from accblas_b200 import _ffi
    L = _ffi.lib()
    syms = header_symbols()
    assert len(syms) >= 13
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/accblas_b200.h but not exported"
    assert sorted(_ffi.SYMBOLS) == syms


def test_version_and_strings():
    import accblas_b200 as ab
    assert ab.version() == 100
    L = ab.lib()
    assert L.accblas_status_string(0) == b"ok"
    assert L.accblas_status_string(3) == b"shape mismatch"


def test_argument_validation_needs_no_gpu():
    """Status codes for bad arguments are produced before any CUDA call."""
    import accblas_b200 as ab
    L = ab.lib()
    one = np.array([1.0])
    p = one.ctypes.data_as(ctypes.c_void_p)
    assert L.accblas_gemm(None, 9, 0, 0, 1, 1, 1, 1, p, p, p, p) == 1          # dtype
    assert L.accblas_gemm(None, 1, 7, 0, 1, 1, 1, 1, p, p, p, p) == 2          # op
    assert L.accblas_gemm(None, 1, 0, 0, 2, 3, 4, 2, p, p, p, p) == 3          # inner extents differ
    assert L.accblas_gemm(None, 1, 0, 0, 0, 3, 3, 2, p, p, p, None) == 0       # m == 0: nothing to do
    assert L.accblas_gemm(None, 1, 0, 0, 2, 3, 3, 2, p, p, p, None) == 4       # null C
    assert L.accblas_gemv(None, 1, 0, 2, 3, p, p, p, 2, p) == 3                # |x| wrong
    assert L.accblas_gemv(None, 1, 1, 2, 3, p, p, p, 3, p) == 3
    assert L.accblas_dot(None, 1, 0, -1, p, p, p) == 3
    assert L.accblas_dot(None, 1, 0, 4, p, p, None) == 4
    # D = beta * Cin with an empty inner extent (k == 0) reads Cin like any other product: its pitch and its
    # overlap with C are checked before the dispatch (m = 2, n = 4, k = 0)
    buf = np.zeros(64)
    q = lambda off: ctypes.c_void_p(buf.ctypes.data + 8 * off)                                     # noqa: E731
    ex = L.accblas_gemm_ex
    assert ex(None, 1, 0, 0, 2, 0, 0, 4, p, None, 0, None, 0, p, q(32), 3, q(0), 4) == 3            # ldcin < n
    assert ex(None, 1, 0, 0, 2, 0, 0, 4, p, None, 0, None, 0, p, q(2), 4, q(0), 4) == 8             # Cin overlaps C
    assert ex(None, 1, 0, 0, 2, 0, 0, 4, p, None, 0, None, 0, p, q(0), 5, q(0), 4) == 8             # same base, other pitch
    assert ex(None, 1, 0, 0, 2, 0, 0, 4, p, None, 0, None, 0, p, None, 4, q(0), 4) == 4             # beta without Cin
    # strided batch: negative stride / batch, outputs that would overlap, operand stride inside one operand
    sb = L.accblas_gemm_strided_batched
    assert sb(None, 1, 0, 0, 2, 3, 3, 2, p, p, -1, p, 6, p, 4, 2) == 3
    assert sb(None, 1, 0, 0, 2, 3, 3, 2, p, p, 6, p, 6, p, 4, -2) == 3
    assert sb(None, 1, 0, 0, 2, 3, 3, 2, p, p, 6, p, 6, p, 3, 2) == 8           # strideC < m * n
    assert sb(None, 1, 0, 0, 2, 3, 3, 2, p, p, 5, p, 6, p, 4, 2) == 3           # 0 < strideA < rowsA * colsA
    assert sb(None, 1, 0, 0, 2, 3, 4, 2, p, p, 6, p, 8, p, 4, 2) == 3           # inner extents differ
    assert sb(None, 1, 0, 0, 2, 3, 3, 2, p, p, 6, p, 6, p, 4, 0) == 0           # empty batch


def test_kernel_selection_names():
    import accblas_b200 as ab
    assert ab.kernel_name(1, "N", "N", (3, 5), (5, 2)) == "gemm_simt_generic"
    assert ab.kernel_name(1, "N", "N", (3, 5), (4, 2)) == "invalid"
    assert ab.kernel_name(1, "N", "N", (0, 5), (5, 2)) == "none"
    # dispatch policy (csrc/api.cu pick_family), host logic only: families by dtype and size
    sq = lambda n: ((n, n), (n, n))  # noqa: E731
    assert ab.kernel_name(1, "N", "T", *sq(8192)) == "dgemm_dmma_tma"
    assert ab.kernel_name(0, "N", "N", *sq(8192)) == "sgemm_tf32x3_tcgen05"
    assert ab.kernel_name(0, "N", "N", *sq(1000)) == "sgemm_tf32x3_tcgen05"      # 128 x 64 tcgen05 tile
    assert ab.kernel_name(0, "N", "N", *sq(500)) == "sgemm_tf32x3_mma"
    assert ab.kernel_name(0, "N", "N", *sq(250)) == "sgemm_tf32x3_mma_direct"    # 1000-byte rows: the kernel loads its panels itself
    assert ab.kernel_name(0, "N", "N", (250, 802), (802, 250)) in ("sgemm_tf32x3_mma", "sgemm_tf32x3_tcgen05")   # k > 768: via the re-pitch copy
    assert ab.kernel_name(1, "N", "N", *sq(251)) in ("dgemm_dmma_direct", "dgemm_dmma_tma")  # (direct needs the SM count: GPU box)
    assert ab.kernel_name(3, "H", "N", *sq(4096)) == "zgemm_3m_dmma"
    assert ab.kernel_name(3, "H", "N", *sq(1000)) == "zgemm_dmma_tma"
    assert ab.kernel_name(2, "H", "N", *sq(4096)) == "cgemm_3m_tf32x3_tcgen05"
    assert ab.kernel_name(2, "H", "N", *sq(1000)) == "cgemm_tf32x3_tcgen05"
    assert ab.kernel_name(2, "N", "N", *sq(250)) == "cgemm_tf32x3_mma"
    assert ab.kernel_name(1, "N", "N", (4096, 1), (1, 4096)) == "gemm_rank1_outer"
    try:
        ab.set_tuning(ab.TUNE_ZGEMM_3M, 0)
        ab.set_tuning(ab.TUNE_CGEMM_3M, 0)
        assert ab.kernel_name(3, "H", "N", *sq(4096)) == "zgemm_dmma_tma"
        assert ab.kernel_name(2, "H", "N", *sq(4096)) == "cgemm_tf32x3_tcgen05"
    finally:
        ab.set_tuning(ab.TUNE_ZGEMM_3M, -1)
        ab.set_tuning(ab.TUNE_CGEMM_3M, -1)


def test_python_mirror_rejects_bad_inputs():
    import accblas_b200 as ab
    with pytest.raises(ValueError):
        ab.gemm(1, "X", np.zeros((2, 2)), "N", np.zeros((2, 2)))
    with pytest.raises(TypeError):
        ab.gemm(1, "N", np.zeros((2, 2), dtype=np.int32), "N", np.zeros((2, 2), dtype=np.int32))
    with pytest.raises(TypeError):
        ab.dotc(np.zeros(3), np.zeros(3))
    with pytest.raises(TypeError):
        ab.gemm(1, "N", np.zeros((2, 2)), "N", np.zeros((2, 2), dtype=np.float32))


def test_no_cpu_fallback_in_product():
    """The product tree must not import / call the oracle, a CPU BLAS, torch matmul or cuBLAS.
    Python: imports and identifiers are checked on the AST (docstrings may cite the reference);
    CUDA/C++: code with comments stripped."""
    import ast
    bad = re.compile(r"oracle|openblas|scipy|cublas|matmul|\bmm\b|einsum|\bdot\b(?=\()", re.I)
    prod = os.path.join(ROOT, "accelerate-blas_b200")
    seen = 0
    for d, _, files in os.walk(prod):
        if os.sep + "build" in d or os.sep + "haskell" in d:
            continue
        for f in files:
            path = os.path.join(d, f)
            if f.endswith(".py"):
                seen += 1
                tree = ast.parse(open(path).read())
                for node in ast.walk(tree):
                    names = []
                    if isinstance(node, ast.Import):
                        names = [a.name for a in node.names]
                    elif isinstance(node, ast.ImportFrom):
                        names = [node.module or ""] + [a.name for a in node.names]
                    elif isinstance(node, ast.Attribute):
                        names = [node.attr] if node.attr in ("matmul", "mm", "bmm", "einsum", "mv", "addmm") else []
                    for nm in names:
                        assert not re.search(r"oracle|openblas|scipy|cublas|matmul|einsum|^mm$|^bmm$|^mv$|^addmm$", nm, re.I), \
                            f"{path}: {nm}"
            elif f.endswith((".cu", ".cuh", ".cpp", ".h")):
                seen += 1
                src = open(path, errors="ignore").read()
                src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
                src = re.sub(r"//[^\n]*", "", src)
                m = re.search(r"oracle|openblas|\bcblas_|cublas", src, re.I)
                assert not m, f"{path}: {m.group(0)}"
    assert seen >= 8


def test_cpp_mirror_builds_and_links():
    """include/accblas_b200.hpp (the C++ host-side mirror of the reference's operator interface) compiles with a
    plain g++ and links against the library; running it needs a GPU (tests/test_parity_gpu.py)."""
    import subprocess
    cpp = os.path.join(ROOT, "tests", "cpp")
    subprocess.check_call(["make", "-s", "-C", cpp])
    out = subprocess.run(["ldd", os.path.join(cpp, "hpp_check")], capture_output=True, text=True).stdout
    assert "libaccblas_b200.so" in out and "not found" not in out.split("libaccblas_b200.so")[1].splitlines()[0]
