"""tools/exp/host_trace.py <tree> [raise_default_pool] -- timeline of one accblas_gemm_host dgemm 8192^3 (ACCBLAS_HOST_TRACE=1:
ms since the call's start at which every panel upload / strip product finished), for the library of <tree>."""
import ctypes
import json
import os
import sys
import time

tree = os.path.abspath(sys.argv[1])
sys.path.insert(0, os.path.join(tree, "accelerate-blas_b200", "python"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import accblas_b200 as ab  # noqa: E402

n = 8192
L = ab.lib()
if len(sys.argv) > 2:          # r1 behaviour of the default pool: never give freed blocks back
    rt = ctypes.CDLL("libcudart.so.12")
    pool = ctypes.c_void_p()
    assert rt.cudaDeviceGetDefaultMemPool(ctypes.byref(pool), 0) == 0
    keep = ctypes.c_uint64(2 ** 64 - 1)
    assert rt.cudaMemPoolSetAttribute(pool, 4, ctypes.byref(keep)) == 0      # cudaMemPoolAttrReleaseThreshold = 4
Ah = torch.rand((n, n), dtype=torch.float64).pin_memory()
Bh = torch.rand((n, n), dtype=torch.float64).pin_memory()
Ch = torch.empty((n, n), dtype=torch.float64).pin_memory()
one = np.array([1.0])
vp = ctypes.c_void_p
ts = []
for i in range(6):
    if i == 5:
        os.environ["ACCBLAS_HOST_TRACE"] = "1"
    torch.cuda.synchronize()
    t = time.perf_counter()
    assert L.accblas_gemm_host(None, 1, 0, 0, n, n, n, n, one.ctypes.data_as(vp), vp(Ah.data_ptr()), vp(Bh.data_ptr()), vp(Ch.data_ptr())) == 0
    ts.append(round((time.perf_counter() - t) * 1e3, 3))
print(json.dumps({"tree": tree, "default_pool_raised": len(sys.argv) > 2, "env_default_pool": os.environ.get("ACCBLAS_DEFAULT_POOL"), "host_ms": ts}), flush=True)
