"""experiment: accblas_gemm_host dgemm 8192^3 from pinned memory vs wavefront panel count"""
import sys, os, time, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
import numpy as np, torch
import accblas_b200 as ab
n = 8192
L = ab.lib()
A = torch.rand(n, n, dtype=torch.float64).pin_memory(); B = torch.rand(n, n, dtype=torch.float64).pin_memory()
C = torch.empty(n, n, dtype=torch.float64).pin_memory()
one = np.array([1.0])
def run():
    st = L.accblas_gemm_host(None, 1, 0, 0, n, n, n, n, one.ctypes.data_as(ctypes.c_void_p), ctypes.c_void_p(A.data_ptr()),
                             ctypes.c_void_p(B.data_ptr()), ctypes.c_void_p(C.data_ptr()))
    assert st == 0
for p in [int(a) for a in sys.argv[1:]] or [16]:
    ab.set_tuning(6, p)
    run(); run()
    ts = []
    for _ in range(4):
        t = time.perf_counter(); run(); ts.append((time.perf_counter() - t) * 1e3)
    print("panels", p, "ms", ["%.2f" % t for t in sorted(ts)], flush=True)
