"""ctypes binding of libaccblas_b200.so (include/accblas_b200.h).

The shared object is the product; this module only declares its symbols.  There is no
fallback: if the library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
PKG_ROOT = os.path.dirname(os.path.dirname(_HERE))               # accelerate-blas_b200/
LIB_PATH = os.environ.get("ACCBLAS_B200_LIB", os.path.join(PKG_ROOT, "lib", "libaccblas_b200.so"))

# every symbol include/accblas_b200.h declares (tests/test_abi.py checks this list against the header)
SYMBOLS = (
    "accblas_gemm", "accblas_gemm_ld", "accblas_gemm_ex", "accblas_last_cuda_error_r", "accblas_launch_count_total",
    "accblas_trim_workspace", "accblas_workspace_cached_bytes", "accblas_gemm_strided_batched", "accblas_gemm_alpha_dev", "accblas_gemv_alpha_dev", "accblas_gemm_workspace_bytes", "accblas_gemv", "accblas_dot", "accblas_dot_compensated", "accblas_dot_static",
    "accblas_gemm_host", "accblas_gemv_host", "accblas_dot_host",
    "accblas_sdot", "accblas_asum", "accblas_amax", "accblas_amin",
    "accblas_sdot_host", "accblas_asum_host", "accblas_amax_host", "accblas_amin_host",
    "accblas_version", "accblas_status_string", "accblas_last_cuda_error",
    "accblas_gemm_kernel_name", "accblas_launch_count", "accblas_set_gemm_path", "accblas_set_tuning",
    "accblas_peer_export", "accblas_peer_open", "accblas_peer_close_all", "accblas_peer_copy2d",
    "accblas_mg_unique_id", "accblas_mg_init", "accblas_mg_finalize", "accblas_mg_rank", "accblas_mg_world",
    "accblas_mg_row_panel", "accblas_mg_k_chunks", "accblas_mg_gemm", "accblas_mg_gemm_plan", "accblas_mg_gemm_run",
    "accblas_mg_gemm_plan_destroy", "accblas_mg_gemv", "accblas_mg_dot",
)


class PeerHandle(ctypes.Structure):
    """accblas_peer_handle_t"""
    _fields_ = [("ipc", ctypes.c_ubyte * 64), ("offset", ctypes.c_uint64), ("size", ctypes.c_uint64),
                ("device", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class AccblasError(RuntimeError):
    def __init__(self, status: int, where: str, detail: str = ""):
        self.status = status
        msg = f"{where}: status {status}"
        if detail:
            msg += f" ({detail})"
        super().__init__(msg)


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OSError(
            f"{LIB_PATH} not found: build it with `make -C '{PKG_ROOT}'` (or __graft_entry__.build()). "
            "accblas_b200 has no CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    i64, vp, ci, sz = ctypes.c_int64, ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
    L.accblas_gemm.argtypes = [vp, ci, ci, ci, i64, i64, i64, i64, vp, vp, vp, vp]
    L.accblas_gemm_host.argtypes = L.accblas_gemm.argtypes
    L.accblas_gemm_ld.argtypes = [vp, ci, ci, ci, i64, i64, i64, i64, vp, vp, i64, vp, i64, vp, i64]
    L.accblas_gemm_ld.restype = ci
    L.accblas_gemm_ex.argtypes = [vp, ci, ci, ci, i64, i64, i64, i64, vp, vp, i64, vp, i64, vp, vp, i64, vp, i64]
    L.accblas_gemm_ex.restype = ci
    L.accblas_last_cuda_error_r.argtypes = [ctypes.c_char_p, sz]
    L.accblas_last_cuda_error_r.restype = sz
    L.accblas_launch_count_total.restype = ctypes.c_uint64
    L.accblas_trim_workspace.argtypes = []
    L.accblas_trim_workspace.restype = ci
    L.accblas_workspace_cached_bytes.argtypes = []
    L.accblas_workspace_cached_bytes.restype = sz
    L.accblas_gemm_strided_batched.argtypes = [vp, ci, ci, ci, i64, i64, i64, i64, vp, vp, i64, vp, i64, vp, i64, i64]
    L.accblas_gemm_strided_batched.restype = ci
    L.accblas_gemv.argtypes = [vp, ci, ci, i64, i64, vp, vp, vp, i64, vp]
    L.accblas_gemv_host.argtypes = L.accblas_gemv.argtypes
    L.accblas_gemm_alpha_dev.argtypes = L.accblas_gemm.argtypes
    L.accblas_gemv_alpha_dev.argtypes = L.accblas_gemv.argtypes
    L.accblas_gemm_alpha_dev.restype = L.accblas_gemv_alpha_dev.restype = ci
    L.accblas_dot.argtypes = [vp, ci, ci, i64, vp, vp, vp]
    L.accblas_dot_host.argtypes = L.accblas_dot.argtypes
    L.accblas_dot_compensated.argtypes = L.accblas_dot.argtypes
    L.accblas_dot_compensated.restype = ci
    L.accblas_dot_static.argtypes = L.accblas_dot.argtypes
    L.accblas_dot_static.restype = ci
    L.accblas_sdot.argtypes = [vp, ci, i64, vp, vp, vp, vp]
    L.accblas_sdot_host.argtypes = L.accblas_sdot.argtypes
    for f in (L.accblas_asum, L.accblas_asum_host, L.accblas_amax, L.accblas_amax_host, L.accblas_amin,
              L.accblas_amin_host):
        f.argtypes = [vp, ci, i64, vp, vp]
        f.restype = ci
    L.accblas_sdot.restype = L.accblas_sdot_host.restype = ci
    for f in (L.accblas_gemm, L.accblas_gemm_host, L.accblas_gemv, L.accblas_gemv_host, L.accblas_dot,
              L.accblas_dot_host, L.accblas_version):
        f.restype = ci
    L.accblas_gemm_workspace_bytes.argtypes = [ci, ci, ci, i64, i64, i64, i64]
    L.accblas_gemm_workspace_bytes.restype = sz
    L.accblas_gemm_kernel_name.argtypes = [ci, ci, ci, i64, i64, i64, i64]
    L.accblas_gemm_kernel_name.restype = ctypes.c_char_p
    L.accblas_status_string.argtypes = [ci]
    L.accblas_status_string.restype = ctypes.c_char_p
    L.accblas_last_cuda_error.restype = ctypes.c_char_p
    L.accblas_launch_count.restype = ctypes.c_uint64
    L.accblas_set_gemm_path.argtypes = [ci]
    L.accblas_set_gemm_path.restype = None
    L.accblas_set_tuning.argtypes = [ci, ci]
    L.accblas_set_tuning.restype = None
    L.accblas_peer_export.argtypes = [vp, ctypes.POINTER(PeerHandle)]
    L.accblas_peer_open.argtypes = [ctypes.POINTER(PeerHandle), ctypes.POINTER(vp)]
    L.accblas_peer_close_all.argtypes = []
    L.accblas_peer_copy2d.argtypes = [vp, vp, sz, vp, sz, sz, sz]
    for f in (L.accblas_peer_export, L.accblas_peer_open, L.accblas_peer_close_all, L.accblas_peer_copy2d):
        f.restype = ci
    pi64 = ctypes.POINTER(i64)
    L.accblas_mg_unique_id.argtypes = [vp]
    L.accblas_mg_init.argtypes = [ctypes.POINTER(vp), ci, ci, vp]
    L.accblas_mg_finalize.argtypes = [vp]
    L.accblas_mg_rank.argtypes = [vp]
    L.accblas_mg_world.argtypes = [vp]
    L.accblas_mg_row_panel.argtypes = [i64, ci, ci, pi64, pi64]
    L.accblas_mg_k_chunks.argtypes = [i64, pi64, pi64, ci]
    L.accblas_mg_gemm.argtypes = [vp, vp, ci, ci, ci, i64, i64, i64, i64, vp, vp, vp, vp, ci]
    L.accblas_mg_gemm_plan.argtypes = [vp, ci, ci, ci, i64, i64, i64, i64, vp, vp, vp, ci, ctypes.POINTER(vp)]
    L.accblas_mg_gemm_run.argtypes = [vp, vp, vp]
    L.accblas_mg_gemm_plan_destroy.argtypes = [vp]
    L.accblas_mg_gemv.argtypes = [vp, vp, ci, ci, i64, i64, vp, vp, vp, i64, vp]
    L.accblas_mg_dot.argtypes = [vp, vp, ci, ci, i64, vp, vp, vp]
    for f in (L.accblas_mg_unique_id, L.accblas_mg_init, L.accblas_mg_finalize, L.accblas_mg_rank, L.accblas_mg_world,
              L.accblas_mg_row_panel, L.accblas_mg_k_chunks, L.accblas_mg_gemm, L.accblas_mg_gemm_plan,
              L.accblas_mg_gemm_run, L.accblas_mg_gemm_plan_destroy, L.accblas_mg_gemv, L.accblas_mg_dot):
        f.restype = ci
    _lib = L
    return L


def check(status: int, where: str) -> None:
    if status != 0:
        L = lib()
        detail = L.accblas_status_string(status).decode()
        if status == 5:
            detail += ": " + L.accblas_last_cuda_error().decode()
        raise AccblasError(status, where, detail)
