"""oracle/openblas_ref.py -- TEST INFRASTRUCTURE ONLY (see oracle/oracle.c header).

The reference's llvm-native path restated call for call: its ``native.gemm`` /
``native.gemv`` hooks make exactly one ``cblas_?gemm`` / ``cblas_?gemv`` call with
``CblasRowMajor`` and tight leading dimensions

    C.dgemm C.RowMajor opA' opB' m n k alpha' ptr_A lda ptr_B ldb 0 ptr_C n
        -- src/.../LLVM/Native/Level3.hs:84-87
    C.dgemv C.RowMajor opA' rowsA colsA alpha' ptr_A colsA ptr_x 1 0 ptr_y 1
        -- src/.../LLVM/Native/Level2.hs:67-70

through the Hackage package ``blas-hs-0.1.1.0`` (stack-8.10.yaml:14) into whatever
system CBLAS is linked.  That package and the BLAS are third-party and absent from
/root/reference; here the same CBLAS entry points are reached in the OpenBLAS that
ships inside the scipy wheel (symbols carry a ``scipy_`` prefix).  This module is

  * the second parity pin for the oracle (tests/test_oracle.py), and
  * the timed CPU arm of bench.py (``cpu_baseline`` and ``--impl reference``),
    kind = "port" (same library routine and arguments, not the Haskell binary).

dot: the reference has NO BLAS dot -- ``dotu`` is Accelerate's own fold
(BLAS/Level1.hs:73-74).  ``cblas_?dot*`` is offered here as a labelled stand-in.
"""
from __future__ import annotations

import ctypes
import glob
import os

import numpy as np

CblasRowMajor = 101
CBLAS_OP = {0: 111, 1: 112, 2: 113, "N": 111, "T": 112, "H": 113}  # NoTrans, Trans, ConjTrans

_lib = None


def _find() -> str:
    import scipy

    root = os.path.join(os.path.dirname(os.path.dirname(scipy.__file__)), "scipy.libs")
    cands = sorted(glob.glob(os.path.join(root, "libscipy_openblas*.so")))
    if not cands:
        raise OSError("scipy-bundled OpenBLAS not found under " + root)
    return cands[0]


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(_find())
        _lib.scipy_openblas_get_config.restype = ctypes.c_char_p
        _lib.scipy_openblas_get_num_threads.restype = ctypes.c_int
        _lib.scipy_cblas_sdot.restype = ctypes.c_float
        _lib.scipy_cblas_ddot.restype = ctypes.c_double
    return _lib


def config() -> str:
    return lib().scipy_openblas_get_config().decode()


def set_threads(n: int) -> int:
    lib().scipy_openblas_set_num_threads(ctypes.c_int(int(n)))
    return lib().scipy_openblas_get_num_threads()


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:  # pragma: no cover
        return os.cpu_count() or 1


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _scalar(alpha, dtype):
    """real alpha by value, complex alpha by pointer (CBLAS convention)."""
    dtype = np.dtype(dtype)
    if dtype == np.float32:
        return ctypes.c_float(float(alpha)), None
    if dtype == np.float64:
        return ctypes.c_double(float(alpha)), None
    buf = np.array([alpha], dtype=dtype)
    return _p(buf), buf


_PFX = {np.dtype(np.float32): "s", np.dtype(np.float64): "d",
        np.dtype(np.complex64): "c", np.dtype(np.complex128): "z"}


def gemm(alpha, opA, A: np.ndarray, opB, B: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
    """One cblas_?gemm(RowMajor, ...) call with the reference's arguments (Native/Level3.hs:84-87)."""
    assert A.flags.c_contiguous and B.flags.c_contiguous and A.dtype == B.dtype
    oa, ob = CBLAS_OP[opA], CBLAS_OP[opB]
    rowsA, colsA = A.shape
    rowsB, colsB = B.shape
    m, k = (rowsA, colsA) if oa == 111 else (colsA, rowsA)    # Native/Level3.hs:56-58
    n = colsB if ob == 111 else rowsB                         # Native/Level3.hs:59-61
    kb = rowsB if ob == 111 else colsB
    if k != kb:
        raise ValueError("inner extents differ")
    C = out if out is not None else np.empty((m, n), dtype=A.dtype)
    al, keep_a = _scalar(alpha, A.dtype)
    be, keep_b = _scalar(0, A.dtype)
    fn = getattr(lib(), f"scipy_cblas_{_PFX[A.dtype]}gemm")
    ci = ctypes.c_int
    fn(ci(CblasRowMajor), ci(oa), ci(ob), ci(m), ci(n), ci(k), al, _p(A), ci(colsA), _p(B), ci(colsB),
       be, _p(C), ci(n))
    return C


def gemv(alpha, opA, A: np.ndarray, x: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
    """One cblas_?gemv(RowMajor, ...) call with the reference's arguments (Native/Level2.hs:67-70)."""
    assert A.flags.c_contiguous and x.flags.c_contiguous and A.dtype == x.dtype
    oa = CBLAS_OP[opA]
    rowsA, colsA = A.shape
    leny = rowsA if oa == 111 else colsA                      # Native/Level2.hs:46-48
    if x.shape[0] != (colsA if oa == 111 else rowsA):
        raise ValueError("vector extent differs")
    y = out if out is not None else np.empty((leny,), dtype=A.dtype)
    al, keep_a = _scalar(alpha, A.dtype)
    be, keep_b = _scalar(0, A.dtype)
    fn = getattr(lib(), f"scipy_cblas_{_PFX[A.dtype]}gemv")
    ci = ctypes.c_int
    fn(ci(CblasRowMajor), ci(oa), ci(rowsA), ci(colsA), al, _p(A), ci(colsA), _p(x), ci(1), be, _p(y), ci(1))
    return y


def dot(x: np.ndarray, y: np.ndarray, conj: bool = False):
    """cblas_?dot / ?dotu_sub / ?dotc_sub -- STAND-IN (the reference has no BLAS dot)."""
    assert x.flags.c_contiguous and y.flags.c_contiguous and x.dtype == y.dtype
    n = min(x.shape[0], y.shape[0])
    ci = ctypes.c_int
    L = lib()
    if x.dtype == np.float32:
        return np.float32(L.scipy_cblas_sdot(ci(n), _p(x), ci(1), _p(y), ci(1)))
    if x.dtype == np.float64:
        return np.float64(L.scipy_cblas_ddot(ci(n), _p(x), ci(1), _p(y), ci(1)))
    out = np.zeros((1,), dtype=x.dtype)
    name = f"scipy_cblas_{_PFX[x.dtype]}dot{'c' if conj else 'u'}_sub"
    getattr(L, name)(ci(n), _p(x), ci(1), _p(y), ci(1), _p(out))
    return out[0]
