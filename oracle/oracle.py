"""oracle/oracle.py -- TEST INFRASTRUCTURE ONLY (see oracle/oracle.c header).

ctypes loader for ``oracle/liboracle.so`` (the plain-C restatement of the reference's
``mXm`` / ``mXv`` / ``dotu`` / ``dotc`` fallbacks) plus the error metrics the parity
tests use.  Nothing under ``accelerate-blas_b200/`` may import this module; only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / --impl
reference legs do.

Dtype / op codes follow ``Type.hs:32-36`` (NumericR) and ``Type.hs:78-90`` (Transpose)
of the reference and are identical to the C ABI's (include/accblas_b200.h).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")

F32, F64, C32, C64 = 0, 1, 2, 3
OP_N, OP_T, OP_H = 0, 1, 2
OPS = {"N": OP_N, "T": OP_T, "H": OP_H}

NP_DTYPE = {F32: np.float32, F64: np.float64, C32: np.complex64, C64: np.complex128}
DTYPE_OF = {np.dtype(v): k for k, v in NP_DTYPE.items()}
EPS = {F32: 2.0 ** -24, F64: 2.0 ** -53, C32: 2.0 ** -24, C64: 2.0 ** -53}
NAMES = {F32: "f32", F64: "f64", C32: "c32", C64: "c64"}


def build(force: bool = False) -> str:
    """Compile oracle.c with gcc (oracle/Makefile).  Building the checker is not using it."""
    src = os.path.join(HERE, "oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", HERE, "liboracle.so"])
    return LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(LIB_PATH)
        i64, vp, ci = ctypes.c_int64, ctypes.c_void_p, ctypes.c_int
        L.oracle_gemm.argtypes = [ci, ci, ci, ci, i64, i64, i64, i64, vp, vp, vp, vp]
        L.oracle_gemv.argtypes = [ci, ci, ci, i64, i64, vp, vp, vp, i64, vp]
        L.oracle_dot.argtypes = [ci, ci, ci, i64, vp, vp, vp]
        L.oracle_sdot.argtypes = [ci, ci, i64, vp, vp, vp, vp]
        L.oracle_asum.argtypes = [ci, ci, i64, vp, vp]
        L.oracle_iamax.argtypes = [ci, ci, i64, vp, ctypes.POINTER(ctypes.c_int64)]
        for f in (L.oracle_gemm, L.oracle_gemv, L.oracle_dot, L.oracle_sdot, L.oracle_asum, L.oracle_iamax):
            f.restype = ci
        _lib = L
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def _op(op) -> int:
    return OPS[op] if isinstance(op, str) else int(op)


def _code(a: np.ndarray) -> int:
    return DTYPE_OF[a.dtype]


def gemm(alpha, opA, A: np.ndarray, opB, B: np.ndarray, wide: bool = True) -> np.ndarray:
    """alpha * op(A) * op(B)  --  BLAS/Level3.hs:57-110 (mXm)."""
    A = np.ascontiguousarray(A)
    B = np.ascontiguousarray(B, dtype=A.dtype)
    dt = _code(A)
    oa, ob = _op(opA), _op(opB)
    m = A.shape[0] if oa == OP_N else A.shape[1]
    n = B.shape[1] if ob == OP_N else B.shape[0]
    C = np.empty((m, n), dtype=A.dtype)
    al = np.array([alpha], dtype=A.dtype)
    rc = lib().oracle_gemm(dt, int(wide), oa, ob, A.shape[0], A.shape[1], B.shape[0], B.shape[1],
                           _p(al), _p(A), _p(B), _p(C))
    if rc:
        raise ValueError("oracle_gemm: bad arguments (shape mismatch?)")
    return C


def gemv(alpha, opA, A: np.ndarray, x: np.ndarray, wide: bool = True) -> np.ndarray:
    """alpha * op(A) * x  --  BLAS/Level2.hs:57-93 (mXv)."""
    A = np.ascontiguousarray(A)
    x = np.ascontiguousarray(x, dtype=A.dtype)
    dt = _code(A)
    oa = _op(opA)
    leny = A.shape[0] if oa == OP_N else A.shape[1]
    y = np.empty((leny,), dtype=A.dtype)
    al = np.array([alpha], dtype=A.dtype)
    rc = lib().oracle_gemv(dt, int(wide), oa, A.shape[0], A.shape[1], _p(al), _p(A), _p(x), x.shape[0], _p(y))
    if rc:
        raise ValueError("oracle_gemv: bad arguments (shape mismatch?)")
    return y


def dot(x: np.ndarray, y: np.ndarray, conj: bool = False, wide: bool = True):
    """dotu / dotc -- BLAS/Level1.hs:73-74, 85-92; length = min(|x|,|y|) (zipWith)."""
    x = np.ascontiguousarray(x)
    y = np.ascontiguousarray(y, dtype=x.dtype)
    n = min(x.shape[0], y.shape[0])
    out = np.zeros((1,), dtype=x.dtype)
    rc = lib().oracle_dot(_code(x), int(wide), int(conj), n, _p(x), _p(y), _p(out))
    if rc:
        raise ValueError("oracle_dot: bad arguments")
    return out[0]


def kbn_sum(xs):
    """`sum kbn xs` -- Numeric/Sum.hs:81-84: `map from . fold add zero . map into` with the Kahan-Babuska-Neumaier
    state `KBN s c`, `into x = KBN x 0`, `from (KBN s c) = s + c` (:128-132) and `add = kbnAdd` (:119-125):
        s' = s1 + s2;  c' = c1 + c2 + (if |s1| >= |s2| then (s1 - s') + s2 else (s2 - s') + s1)
    Restated as the sequential left fold in the element type (pure Python loop: small cases only)."""
    xs = np.asarray(xs)
    ty = xs.dtype.type
    s, c = ty(0), ty(0)
    for v in xs:
        s2, c2 = ty(v), ty(0)
        sn = ty(s + s2)
        lost = ty(ty(s - sn) + s2) if abs(s) >= abs(s2) else ty(ty(s2 - sn) + s)
        c = ty(ty(c + c2) + lost)
        s = sn
    return ty(s + c)


def dot_exact(x: np.ndarray, y: np.ndarray, conj: bool = False):
    """The dot product in exact rational arithmetic (fractions.Fraction), rounded once to double / complex double:
    the truth the compensated dot (accblas_dot_compensated) is measured against.  O(n) big-integer work: n <~ 10^5."""
    from fractions import Fraction as Fr
    x = np.asarray(x)
    y = np.asarray(y)
    n = min(x.shape[0], y.shape[0])
    if np.iscomplexobj(x):
        sg = -1 if conj else 1
        re, im = Fr(0), Fr(0)
        for a, b in zip(x[:n], y[:n]):
            ar, ai, br, bi = Fr(float(a.real)), sg * Fr(float(a.imag)), Fr(float(b.real)), Fr(float(b.imag))
            re += ar * br - ai * bi
            im += ar * bi + ai * br
        return complex(float(re), float(im))
    acc = Fr(0)
    for a, b in zip(x[:n], y[:n]):
        acc += Fr(float(a)) * Fr(float(b))
    return float(acc)


def sdot(z, x: np.ndarray, y: np.ndarray, wide: bool = True):
    """sdot z xs ys -- BLAS/Level1.hs:44-62: z + sum x_i y_i in double precision, rounded once."""
    x = np.ascontiguousarray(x)
    y = np.ascontiguousarray(y, dtype=x.dtype)
    n = min(x.shape[0], y.shape[0])
    zz = np.array([z], dtype=x.dtype)
    out = np.zeros((1,), dtype=x.dtype)
    if lib().oracle_sdot(_code(x), int(wide), n, _p(zz), _p(x), _p(y), _p(out)):
        raise ValueError("oracle_sdot: bad arguments")
    return out[0]


def asum(x: np.ndarray, wide: bool = True):
    """asum -- BLAS/Level1.hs:100-109; result in the element's base type."""
    x = np.ascontiguousarray(x)
    base = np.float32 if _code(x) in (F32, C32) else np.float64
    out = np.zeros((1,), dtype=base)
    if lib().oracle_asum(_code(x), int(wide), x.shape[0], _p(x), _p(out)):
        raise ValueError("oracle_asum: bad arguments")
    return out[0]


def _iamax(x: np.ndarray, want_max: bool) -> int:
    x = np.ascontiguousarray(x)
    out = ctypes.c_int64(-1)
    if lib().oracle_iamax(_code(x), int(want_max), x.shape[0], _p(x), ctypes.byref(out)):
        raise ValueError("oracle_iamax: empty vector (fold1) or bad dtype")
    return int(out.value)


def amax(x: np.ndarray) -> int:
    """amax -- BLAS/Level1.hs:115-126 (last index among equal maxima)."""
    return _iamax(x, True)


def amin(x: np.ndarray) -> int:
    """amin -- BLAS/Level1.hs:132-143 (last index among equal minima)."""
    return _iamax(x, False)


# ---------------------------------------------------------------------------------
# error metrics: normwise, in units of (k+2) * eps * ||A||_F * ||B||_F
# ---------------------------------------------------------------------------------
def fro(a) -> float:
    a = np.asarray(a)
    return float(np.sqrt(np.sum(np.abs(a.astype(np.complex128 if np.iscomplexobj(a) else np.float64)) ** 2)))


def gemm_error_units(C, C_ref, A, B, k: int, alpha=1.0) -> float:
    """||C - C_ref||_F / ((k + 2) * eps * |alpha| * ||A||_F * ||B||_F); eps of C's dtype.

    k roundings of the accumulation plus the two of the alpha scaling and the final store
    (the classical gamma_{k+2} bound); for large k this is north_star's k * eps * ||A|| ||B||."""
    dt = DTYPE_OF[np.asarray(C).dtype]
    denom = (max(k, 0) + 2) * EPS[dt] * max(abs(alpha), 1e-300) * fro(A) * fro(B)
    num = fro(np.asarray(C).astype(np.complex128) - np.asarray(C_ref).astype(np.complex128))
    if denom == 0.0:
        return 0.0 if num == 0.0 else float("inf")
    return num / denom


def similar(u, v, abs_tol=5e-5, rel_tol=5e-3) -> bool:
    """The reference test-suite's `~~~` (Data.Array.Accelerate.Test.Similar, external;
    recalled constants): |u-v| < 5e-5 or |u-v|/max(|u|,|v|) < 5e-3,
    complex parts compared independently, NaN~NaN, Inf~Inf."""
    u = np.asarray(u)
    v = np.asarray(v)
    if np.iscomplexobj(u) or np.iscomplexobj(v):
        return similar(u.real, v.real, abs_tol, rel_tol) and similar(u.imag, v.imag, abs_tol, rel_tol)
    u = u.astype(np.float64)
    v = v.astype(np.float64)
    d = np.abs(u - v)
    mx = np.maximum(np.abs(u), np.abs(v))
    with np.errstate(invalid="ignore", divide="ignore"):
        ok = (d < abs_tol) | (d / mx < rel_tol) | (np.isnan(u) & np.isnan(v)) | (np.isinf(u) & (u == v))
    return bool(np.all(ok))


def uniform(rng: np.random.Generator, shape, dt: int) -> np.ndarray:
    """i.i.d. uniform(-1,1), complex = independent parts -- test/Test/Util.hs:34-38."""
    base = np.float32 if dt in (F32, C32) else np.float64
    if dt in (C32, C64):
        re = rng.uniform(-1, 1, size=shape).astype(base)
        im = rng.uniform(-1, 1, size=shape).astype(base)
        return (re + 1j * im).astype(NP_DTYPE[dt])
    return rng.uniform(-1, 1, size=shape).astype(base)
