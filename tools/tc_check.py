"""tools/tc_check.py -- first-light check of the tcgen05 path against the oracle (small + medium)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
import numpy as np, torch
import accblas_b200 as ab
from oracle import oracle as O
rng = np.random.default_rng(1)
dev = torch.device("cuda:0")
for dt in (O.F32, O.C32):
    for (oa, ob, m, n, k) in [("N", "T", 128, 128, 1024), ("N", "N", 256, 384, 320), ("T", "N", 300, 200, 517), ("H", "H", 257, 129, 1000),
                              ("N", "H", 512, 512, 2048)]:
        A = O.uniform(rng, (m, k) if oa == "N" else (k, m), dt)
        B = O.uniform(rng, (k, n) if ob == "N" else (n, k), dt)
        al = O.uniform(rng, (1,), dt)[0]
        name = ab.kernel_name(dt, oa, ob, A.shape, B.shape)
        C = ab.gemm(al, oa, torch.from_numpy(A).to(dev), ob, torch.from_numpy(B).to(dev)).cpu().numpy()
        ref = O.gemm(al, oa, A, ob, B)
        lit = O.gemm(al, oa, A, ob, B, wide=False)
        print(O.NAMES[dt], oa, ob, m, n, k, name, "err_units", round(O.gemm_error_units(C, ref, A, B, k, al), 4),
              "fp32-literal err_units", round(O.gemm_error_units(lit, ref, A, B, k, al), 4),
              "max abs", float(np.abs(C - ref).max()), flush=True)
