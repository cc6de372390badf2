"""experiment: dgemm 8192 time vs raster group (DRAM traffic is read with ncu around this script)"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
import torch
import accblas_b200 as ab
n = 8192
A = torch.rand(n, n, dtype=torch.float64, device="cuda"); B = torch.rand(n, n, dtype=torch.float64, device="cuda")
for grp in [int(a) for a in sys.argv[1:]] or [16]:
    ab.set_tuning(5, grp)
    for _ in range(2): ab.mXm(A, B)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): ab.mXm(A, B)
    e1.record(); torch.cuda.synchronize()
    print("group", grp, "ms", e0.elapsed_time(e1) / 3, flush=True)
