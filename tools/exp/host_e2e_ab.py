"""tools/exp/host_e2e_ab.py <tree> -- accblas_gemm_host dgemm 8192^3 with the library + python package of <tree> (A/B against
an older checkout)."""
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
Ah = torch.rand((n, n), dtype=torch.float64).pin_memory()
Bh = torch.rand((n, n), dtype=torch.float64).pin_memory()
Ch = torch.empty((n, n), dtype=torch.float64).pin_memory()
one = np.array([1.0])
vp = ctypes.c_void_p
ts = []
for i in range(10):
    torch.cuda.synchronize()
    t = time.perf_counter()
    assert L.accblas_gemm_host(None, 1, 0, 0, n, n, n, n, one.ctypes.data_as(vp), vp(Ah.data_ptr()), vp(Bh.data_ptr()), vp(Ch.data_ptr())) == 0
    ts.append(round((time.perf_counter() - t) * 1e3, 3))
A = Ah.cuda(); B = Bh.cuda()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
ab.gemm(1, "N", A, "N", B); torch.cuda.synchronize()
ev[0].record()
for i in range(5):
    ab.gemm(1, "N", A, "N", B); ev[i + 1].record()
torch.cuda.synchronize()
print(json.dumps({"tree": tree, "host_ms": ts, "device_ms": [round(ev[i].elapsed_time(ev[i + 1]), 3) for i in range(5)]}), flush=True)
