"""accblas_b200 -- host-side mirror of Data.Array.Accelerate.Numeric.LinearAlgebra for the
B200-native library libaccblas_b200.so.

The reference's public surface for the hot path (same names, argument order and meaning):

    gemm alpha opA A opB B        BLAS/Level3.hs:57-64        ->  gemm(alpha, opA, A, opB, B)
    gemv alpha opA A x            BLAS/Level2.hs:57-63        ->  gemv(alpha, opA, A, x)
    dotu xs ys / dotc xs ys       BLAS/Level1.hs:73-74,85-92  ->  dotu(x, y) / dotc(x, y)
    (<>) (#>) (<#) (<.>) (><)     LinearAlgebra.hs:58-155     ->  mXm, mXv, vXm, dot, outer
    Transpose = N | T | H         Type.hs:78-90               ->  N, T, H
    Numeric e => Float, Double, Complex Float, Complex Double (Type.hs:32-51)
                                  ->  float32, float64, complex64, complex128

Arrays are row-major and dense like Accelerate's.  Two kinds of operand are accepted:

  * ``torch.Tensor`` on a CUDA device: the device-pointer ABI (accblas_gemm / gemv / dot) is
    called on torch's current stream -- the analogue of the ForeignAcc hook receiving the
    backend's device pointers and stream (LLVM/PTX/Level3.hs:78-85).  Results are new device
    tensors; ``dotu`` returns a 0-d device tensor (the reference's ``Acc (Scalar e)``).
  * ``numpy.ndarray``: the host-buffer ABI (accblas_*_host) -- what ``run`` does end to end:
    upload, hook, download (test/Test/BLAS/Level3.hs:52-54).

PyTorch is only the provider of device memory and streams.  No arithmetic happens in
Python, torch or numpy, and there is no CPU fallback: without the CUDA library every call
raises.
"""
from __future__ import annotations

import ctypes
from typing import Union

import numpy as np

from . import _ffi
from ._ffi import AccblasError, lib  # noqa: F401

__all__ = ["N", "T", "H", "gemm", "gemm_batched", "gemv", "dotu", "dotc", "dot_compensated", "sdot", "asum", "amax", "amin", "mXm", "mXv", "vXm", "dot", "outer",
           "AccblasError", "kernel_name", "launch_count", "set_gemm_path", "set_tuning", "version"]

# Transpose -- Type.hs:78-90
N, T, H = 0, 1, 2
_OPS = {"N": N, "T": T, "H": H, N: N, T: T, H: H}

# NumericR -- Type.hs:32-36
F32, F64, C32, C64 = 0, 1, 2, 3
_NP_CODE = {np.dtype(np.float32): F32, np.dtype(np.float64): F64,
            np.dtype(np.complex64): C32, np.dtype(np.complex128): C64}
_NP_OF = {v: k for k, v in _NP_CODE.items()}


def _torch():
    import torch
    return torch


def _is_torch(a) -> bool:
    return type(a).__module__.startswith("torch")


def _code(a) -> int:
    if _is_torch(a):
        torch = _torch()
        table = {torch.float32: F32, torch.float64: F64, torch.complex64: C32, torch.complex128: C64}
        if a.dtype not in table:
            raise TypeError(f"accblas_b200: no Numeric instance for {a.dtype} (Type.hs:38-51)")
        return table[a.dtype]
    dt = np.dtype(a.dtype)
    if dt not in _NP_CODE:
        raise TypeError(f"accblas_b200: no Numeric instance for {dt} (Type.hs:38-51)")
    return _NP_CODE[dt]


def _op(op) -> int:
    try:
        return _OPS[op]
    except (KeyError, TypeError):
        raise ValueError(f"accblas_b200: Transpose must be N, T or H, got {op!r}") from None


def _alpha_buf(alpha, code: int):
    buf = np.array([alpha], dtype=_NP_OF[code])
    return buf, buf.ctypes.data_as(ctypes.c_void_p)


def _alpha_on_device(alpha, code: int) -> bool:
    """alpha given as a one-element CUDA tensor (the reference's device `Scalar e`, BLAS/Level3.hs:64):
    the kernel reads it, no host round trip (accblas_*_alpha_dev)."""
    if not (_is_torch(alpha) and alpha.is_cuda):
        return False
    if alpha.numel() != 1 or _code(alpha) != code:
        raise TypeError("alpha on the device must be a one-element tensor of the operands' element type")
    return True


def _dev_operand(a, code: int, ndim: int, name: str):
    if not _is_torch(a) or not a.is_cuda:
        raise TypeError(f"{name}: expected a CUDA tensor when the first operand is one")
    if _code(a) != code:
        raise TypeError(f"{name}: element type differs from the first operand's")
    if a.dim() != ndim:
        raise ValueError(f"{name}: expected rank {ndim}, got {a.dim()}")
    return a if a.is_contiguous() else a.contiguous()


def _row_strided(a) -> bool:
    """rank-2 view whose rows are dense and whose row pitch is at least the row length: a window of
    a larger row-major matrix, which accblas_gemm_ld takes as (pointer, lda) without a copy."""
    return a.dim() == 2 and a.shape[0] > 0 and a.shape[1] > 0 and a.stride(1) == 1 and a.stride(0) >= a.shape[1]


def _dev_matrix(a, code: int, name: str):
    if _is_torch(a) and a.is_cuda and _code(a) == code and _row_strided(a):
        return a, a.stride(0)
    a = _dev_operand(a, code, 2, name)
    return a, a.shape[1]


def _host_operand(a, code: int, ndim: int, name: str) -> np.ndarray:
    a = np.asarray(a)
    if _code(a) != code:
        raise TypeError(f"{name}: element type differs from the first operand's")
    if a.ndim != ndim:
        raise ValueError(f"{name}: expected rank {ndim}, got {a.ndim}")
    return np.ascontiguousarray(a)


def _stream(t) -> ctypes.c_void_p:
    torch = _torch()
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _ptr(t) -> ctypes.c_void_p:
    return ctypes.c_void_p(t.data_ptr() if _is_torch(t) else t.ctypes.data)


# ------------------------------------------------------------------------------------
# Level 3
# ------------------------------------------------------------------------------------
def gemm(alpha, opA, A, opB, B, out=None, beta=None, Cin=None):
    """``C = alpha * op(A) * op(B)`` -- gemm, BLAS/Level3.hs:57-64.

    shape op(A) = (m, k), shape op(B) = (k, n), shape C = (m, n); stored shapes are pre-op.
    ``out`` (device path only): a contiguous (m, n) tensor to write instead of allocating one.
    ``beta``, ``Cin`` (device path only; not in the reference, whose hook fixes beta = 0): adds ``beta * Cin``
    in the kernel's epilogue (accblas_gemm_ex).  ``Cin`` may be ``out`` itself (accumulate in place).
    """
    oa, ob = _op(opA), _op(opB)
    code = _code(A)
    L = lib()
    if beta is not None and Cin is None:
        raise ValueError("gemm: beta needs Cin")
    if Cin is not None and not (_is_torch(A) and not _alpha_on_device(alpha, code)):
        raise ValueError("gemm: beta / Cin are taken on the device path with a host alpha only")
    if _is_torch(A) and _alpha_on_device(alpha, code):
        torch = _torch()
        A = _dev_operand(A, code, 2, "gemm: A")
        B = _dev_operand(B, code, 2, "gemm: B")
        m = A.shape[0] if oa == N else A.shape[1]
        n = B.shape[1] if ob == N else B.shape[0]
        with torch.cuda.device(A.device):
            C = torch.empty((m, n), dtype=A.dtype, device=A.device) if out is None else out
            st = L.accblas_gemm_alpha_dev(_stream(A), code, oa, ob, A.shape[0], A.shape[1], B.shape[0], B.shape[1],
                                          _ptr(alpha), _ptr(A), _ptr(B), _ptr(C))
        _ffi.check(st, "accblas_gemm_alpha_dev")
        return C
    al_keep, al = _alpha_buf(alpha, code)
    if _is_torch(A):
        torch = _torch()
        A, lda = _dev_matrix(A, code, "gemm: A")
        B, ldb = _dev_matrix(B, code, "gemm: B")
        m = A.shape[0] if oa == N else A.shape[1]          # LLVM/PTX/Level3.hs:57-62
        n = B.shape[1] if ob == N else B.shape[0]
        with torch.cuda.device(A.device):
            if out is None:
                C = torch.empty((m, n), dtype=A.dtype, device=A.device)   # allocateRemote: uninitialised
            else:
                if tuple(out.shape) != (m, n) or out.dtype != A.dtype or not out.is_cuda or \
                        not (out.is_contiguous() or _row_strided(out)):
                    raise ValueError("gemm: out must be a CUDA tensor of shape (m, n) and A's dtype with dense rows")
                C = out
            ldc = C.stride(0) if (m > 0 and n > 0 and not C.is_contiguous()) else n
            if Cin is not None:
                if tuple(Cin.shape) != (m, n) or Cin.dtype != A.dtype or not Cin.is_cuda or \
                        not (Cin.is_contiguous() or _row_strided(Cin)):
                    raise ValueError("gemm: Cin must be a CUDA tensor of shape (m, n) and A's dtype with dense rows")
                ldcin = Cin.stride(0) if (m > 0 and n > 0 and not Cin.is_contiguous()) else n
                be_keep, be = _alpha_buf(1 if beta is None else beta, code)
                st = L.accblas_gemm_ex(_stream(A), code, oa, ob, A.shape[0], A.shape[1], B.shape[0], B.shape[1],
                                       al, _ptr(A), lda, _ptr(B), ldb, be, _ptr(Cin), ldcin, _ptr(C), ldc)
                _ffi.check(st, "accblas_gemm_ex")
                return C
            if lda == A.shape[1] and ldb == B.shape[1] and ldc == n:
                st = L.accblas_gemm(_stream(A), code, oa, ob, A.shape[0], A.shape[1], B.shape[0], B.shape[1],
                                    al, _ptr(A), _ptr(B), _ptr(C))
            else:   # windows of larger matrices: explicit row pitches
                st = L.accblas_gemm_ld(_stream(A), code, oa, ob, A.shape[0], A.shape[1], B.shape[0], B.shape[1],
                                       al, _ptr(A), lda, _ptr(B), ldb, _ptr(C), ldc)
        _ffi.check(st, "accblas_gemm")
        return C
    A = _host_operand(A, code, 2, "gemm: A")
    B = _host_operand(B, code, 2, "gemm: B")
    m = A.shape[0] if oa == N else A.shape[1]
    n = B.shape[1] if ob == N else B.shape[0]
    C = np.empty((m, n), dtype=A.dtype)
    st = L.accblas_gemm_host(None, code, oa, ob, A.shape[0], A.shape[1], B.shape[0], B.shape[1],
                             al, _ptr(A), _ptr(B), _ptr(C))
    _ffi.check(st, "accblas_gemm_host")
    return C


def gemm_batched(alpha, opA, A, opB, B):
    """``C[b] = alpha * op(A[b]) * op(B[b])`` for a stack of products (accblas_gemm_strided_batched).

    A, B: contiguous CUDA tensors of rank 3 (batch, rows, cols) -- or rank 2, which is broadcast over the batch
    (stride 0).  Not an operator of the reference: Accelerate has no batched product.
    """
    oa, ob = _op(opA), _op(opB)
    code = _code(A)
    torch = _torch()
    A = _dev_operand(A, code, A.dim(), "gemm_batched: A")
    B = _dev_operand(B, code, B.dim(), "gemm_batched: B")
    if A.dim() not in (2, 3) or B.dim() not in (2, 3) or (A.dim() == 2 and B.dim() == 2):
        raise ValueError("gemm_batched: operands must be rank 3 (at most one of them rank 2)")
    batch = A.shape[0] if A.dim() == 3 else B.shape[0]
    if A.dim() == 3 and B.dim() == 3 and A.shape[0] != B.shape[0]:
        raise ValueError("gemm_batched: batch extents differ")
    ra, ca = A.shape[-2], A.shape[-1]
    rb, cb = B.shape[-2], B.shape[-1]
    m = ra if oa == N else ca
    n = cb if ob == N else rb
    al_keep, al = _alpha_buf(alpha, code)
    with torch.cuda.device(A.device):
        C = torch.empty((batch, m, n), dtype=A.dtype, device=A.device)
        st = lib().accblas_gemm_strided_batched(_stream(A), code, oa, ob, ra, ca, rb, cb, al,
                                                _ptr(A), ra * ca if A.dim() == 3 else 0,
                                                _ptr(B), rb * cb if B.dim() == 3 else 0, _ptr(C), m * n, batch)
    _ffi.check(st, "accblas_gemm_strided_batched")
    return C


# ------------------------------------------------------------------------------------
# Level 2
# ------------------------------------------------------------------------------------
def gemv(alpha, opA, A, x):
    """``y = alpha * op(A) * x`` -- gemv, BLAS/Level2.hs:57-63."""
    oa = _op(opA)
    code = _code(A)
    L = lib()
    on_dev = _is_torch(A) and _alpha_on_device(alpha, code)
    if not on_dev:
        al_keep, al = _alpha_buf(alpha, code)
    if _is_torch(A):
        torch = _torch()
        A = _dev_operand(A, code, 2, "gemv: A")
        x = _dev_operand(x, code, 1, "gemv: x")
        leny = A.shape[0] if oa == N else A.shape[1]       # LLVM/PTX/Level2.hs:80-82
        with torch.cuda.device(A.device):
            y = torch.empty((leny,), dtype=A.dtype, device=A.device)
            fn = L.accblas_gemv_alpha_dev if on_dev else L.accblas_gemv
            st = fn(_stream(A), code, oa, A.shape[0], A.shape[1], _ptr(alpha) if on_dev else al, _ptr(A), _ptr(x),
                    x.shape[0], _ptr(y))
        _ffi.check(st, "accblas_gemv")
        return y
    A = _host_operand(A, code, 2, "gemv: A")
    x = _host_operand(x, code, 1, "gemv: x")
    leny = A.shape[0] if oa == N else A.shape[1]
    y = np.empty((leny,), dtype=A.dtype)
    st = L.accblas_gemv_host(None, code, oa, A.shape[0], A.shape[1], al, _ptr(A), _ptr(x), x.shape[0], _ptr(y))
    _ffi.check(st, "accblas_gemv_host")
    return y


# ------------------------------------------------------------------------------------
# Level 1
# ------------------------------------------------------------------------------------
def _dot(x, y, conj: bool):
    code = _code(x)
    L = lib()
    if _is_torch(x):
        torch = _torch()
        x = _dev_operand(x, code, 1, "dot: x")
        y = _dev_operand(y, code, 1, "dot: y")
        n = min(x.shape[0], y.shape[0])                    # zipWith intersects extents
        with torch.cuda.device(x.device):
            out = torch.empty((), dtype=x.dtype, device=x.device)
            st = L.accblas_dot(_stream(x), code, int(conj), n, _ptr(x), _ptr(y), _ptr(out))
        _ffi.check(st, "accblas_dot")
        return out
    x = _host_operand(x, code, 1, "dot: x")
    y = _host_operand(y, code, 1, "dot: y")
    n = min(x.shape[0], y.shape[0])
    out = np.zeros((1,), dtype=x.dtype)
    st = L.accblas_dot_host(None, code, int(conj), n, _ptr(x), _ptr(y), _ptr(out))
    _ffi.check(st, "accblas_dot_host")
    return out[0]


def dotu(x, y):
    """``sum_i x_i * y_i`` -- dotu, BLAS/Level1.hs:73-74."""
    return _dot(x, y, False)


def dotc(x, y):
    """``sum_i conj(x_i) * y_i`` -- dotc, BLAS/Level1.hs:85-92 (complex element types only)."""
    if _code(x) not in (C32, C64):
        raise TypeError("dotc: Numeric (Complex e) required (BLAS/Level1.hs:85)")
    return _dot(x, y, True)


def dot_compensated(x, y, conj: bool = False):
    """dotu / dotc (``conj``) with compensated arithmetic: exact product splitting, TwoSum accumulation and the
    reference's Kahan-Babuska-Neumaier combine (Numeric/Sum.hs:119-125) between partial sums -- as accurate as a dot
    evaluated in twice double precision and rounded once.  Device tensors only (result: 0-d device tensor)."""
    code = _code(x)
    if conj and code not in (C32, C64):
        raise TypeError("dot_compensated: conj needs Numeric (Complex e) (BLAS/Level1.hs:85)")
    if not _is_torch(x):
        raise TypeError("dot_compensated: device tensors only")
    torch = _torch()
    x = _dev_operand(x, code, 1, "dot_compensated: x")
    y = _dev_operand(y, code, 1, "dot_compensated: y")
    n = min(x.shape[0], y.shape[0])
    with torch.cuda.device(x.device):
        out = torch.empty((), dtype=x.dtype, device=x.device)
        st = lib().accblas_dot_compensated(_stream(x), code, int(conj), n, _ptr(x), _ptr(y), _ptr(out))
    _ffi.check(st, "accblas_dot_compensated")
    return out


def sdot(z, x, y):
    """``z + sum_i x_i * y_i`` with double-precision products and accumulation, rounded once to
    the element type -- sdot, BLAS/Level1.hs:44-62."""
    code = _code(x)
    z_keep, zp = _alpha_buf(z, code)
    L = lib()
    if _is_torch(x):
        torch = _torch()
        x = _dev_operand(x, code, 1, "sdot: x")
        y = _dev_operand(y, code, 1, "sdot: y")
        n = min(x.shape[0], y.shape[0])
        with torch.cuda.device(x.device):
            out = torch.empty((), dtype=x.dtype, device=x.device)
            st = L.accblas_sdot(_stream(x), code, n, zp, _ptr(x), _ptr(y), _ptr(out))
        _ffi.check(st, "accblas_sdot")
        return out
    x = _host_operand(x, code, 1, "sdot: x")
    y = _host_operand(y, code, 1, "sdot: y")
    out = np.zeros((1,), dtype=x.dtype)
    st = L.accblas_sdot_host(None, code, min(x.shape[0], y.shape[0]), zp, _ptr(x), _ptr(y), _ptr(out))
    _ffi.check(st, "accblas_sdot_host")
    return out[0]


def _reduce1(x, fn_dev, fn_host, out_np_dtype, where):
    code = _code(x)
    L = lib()
    if _is_torch(x):
        torch = _torch()
        x = _dev_operand(x, code, 1, where + ": x")
        tdt = {np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64,
               np.dtype(np.int64): torch.int64}[np.dtype(out_np_dtype(code))]
        with torch.cuda.device(x.device):
            out = torch.empty((), dtype=tdt, device=x.device)
            st = getattr(L, fn_dev)(_stream(x), code, x.shape[0], _ptr(x), _ptr(out))
        _ffi.check(st, fn_dev)
        return out
    x = _host_operand(x, code, 1, where + ": x")
    out = np.zeros((1,), dtype=out_np_dtype(code))
    st = getattr(L, fn_host)(None, code, x.shape[0], _ptr(x), _ptr(out))
    _ffi.check(st, fn_host)
    return out[0]


def _base(code):
    return np.float32 if code in (F32, C32) else np.float64


def asum(x):
    """``sum_i |x_i|`` (complex: ``|re| + |im|``), in the element's base type -- asum, BLAS/Level1.hs:100-109."""
    return _reduce1(x, "accblas_asum", "accblas_asum_host", _base, "asum")


def amax(x):
    """Index (int64) of the element of largest magnitude, the LAST one among ties -- amax,
    BLAS/Level1.hs:115-126 (`cmp ix iy = snd ix > snd iy ? ix : iy`)."""
    return _reduce1(x, "accblas_amax", "accblas_amax_host", lambda c: np.int64, "amax")


def amin(x):
    """Index (int64) of the element of smallest magnitude, the LAST one among ties -- amin,
    BLAS/Level1.hs:132-143."""
    return _reduce1(x, "accblas_amin", "accblas_amin_host", lambda c: np.int64, "amin")


# ------------------------------------------------------------------------------------
# operators -- LinearAlgebra.hs
# ------------------------------------------------------------------------------------
def dot(x, y):
    """``x <.> y = dotu x y`` -- LinearAlgebra.hs:58-60."""
    return dotu(x, y)


def outer(x, y):
    """``x >< y`` : (len x, 1) <> (1, len y) -- LinearAlgebra.hs:73-78."""
    return mXm(x.reshape(-1, 1), y.reshape(1, -1))


def mXv(m, x):
    """``m #> x = gemv 1 N m x`` -- LinearAlgebra.hs:99-101."""
    return gemv(1, N, m, x)


def vXm(x, m):
    """``x <# m = gemv 1 T m x`` -- LinearAlgebra.hs:119-121."""
    return gemv(1, T, m, x)


def mXm(a, b):
    """``a <> b = gemm 1 N a N b`` -- LinearAlgebra.hs:153-155."""
    return gemm(1, N, a, N, b)


# ------------------------------------------------------------------------------------
# introspection
# ------------------------------------------------------------------------------------
def kernel_name(dtype_code: int, opA, opB, shapeA, shapeB) -> str:
    return lib().accblas_gemm_kernel_name(dtype_code, _op(opA), _op(opB), shapeA[0], shapeA[1],
                                          shapeB[0], shapeB[1]).decode()


def workspace_bytes(dtype_code: int, opA, opB, shapeA, shapeB) -> int:
    """Stream-ordered scratch accblas_gemm takes from the library's pool for this problem."""
    return int(lib().accblas_gemm_workspace_bytes(dtype_code, _op(opA), _op(opB), shapeA[0], shapeA[1],
                                                  shapeB[0], shapeB[1]))


def trim_workspace() -> None:
    """Return the library's cached scratch to the driver (accblas_trim_workspace)."""
    _ffi.check(lib().accblas_trim_workspace(), "accblas_trim_workspace")


def launch_count() -> int:
    return int(lib().accblas_launch_count())


def set_gemm_path(path: Union[int, str]) -> None:
    """0/'auto', 1/'generic' (SIMT), 2/'tensor' (error if the problem does not qualify)."""
    table = {"auto": 0, "generic": 1, "tensor": 2}
    lib().accblas_set_gemm_path(table.get(path, path))


TUNE_DMMA_TILE, TUNE_ZMMA_TILE = 0, 1
TUNE_SGEMM_FAMILY, TUNE_SMMA_TILE, TUNE_HOST_PIPELINE, TUNE_RASTER_GROUP, TUNE_HOST_PANELS = 2, 3, 4, 5, 6
TUNE_TC_TILE = 8        # tcgen05 f32/c32 gemm CTA tile: 0 = 128x256, 1 = 128x128, 2 = 128x64
TUNE_CGEMM_3M = 9       # c32 gemm on tcgen05: 0 = 4M real embedding, 1 = 3M whenever possible, -1 = cost model
TUNE_SNAKE_K = 10       # f64 gemm: 0 = every tile walks K forwards
TUNE_STREAMK = 11       # f64 gemm: 0 = never the persistent kernel, 1 = whole-tile rounds + stream-K tail, 2 = pure stream-K
TUNE_RELEASE = 12       # f64/c64 gemm stage release: 0 = fenced everywhere, 2 = delayed by one K-block everywhere
TUNE_MG_SELF_PEER = 14  # accblas_mg_gemm_plan: 1 = the root runs the peers' pull / K-chunk pipeline itself (test hook)
TUNE_DOT_SHAPE = 15      # dot streaming pass CTA shape (see include/accblas_b200.h)
TUNE_MG_DOT_FUSED = 16   # accblas_mg_dot: 0 = three launches (local dot + exchange kernel), default one fused launch
TUNE_MN_ONEBOX = 13     # f64 gemm: 0 = MN-major operand panels as 16 x 16 boxes, default one 3-D box per stage
TUNE_ZGEMM_3M = 7       # 0: complex-double gemm always 4M in registers, 1: 3M whenever possible, -1: cost model


def set_tuning(knob: int, value: int) -> None:
    """accblas_set_tuning: force a kernel variant on this thread (value < 0: automatic)."""
    lib().accblas_set_tuning(int(knob), int(value))


def version() -> int:
    return int(lib().accblas_version())
