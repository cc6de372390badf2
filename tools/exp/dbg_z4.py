import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
import torch
import accblas_b200 as ab
torch.manual_seed(0)
def rnd(*s): return torch.view_as_complex(torch.rand(*s, 2, dtype=torch.float64, device="cuda") * 2 - 1)
for tile in (2, 3, 4, 5):
    ab.set_tuning(1, tile)
    for (m, n, k) in [(768, 768, 96), (2048, 2048, 96), (1024, 1024, 64)]:
        for ops in ("NN", "NT", "TN", "TT"):
            A = rnd(m, k) if ops[0] == "N" else rnd(k, m)
            B = rnd(k, n) if ops[1] == "N" else rnd(n, k)
            ref = (A if ops[0] == "N" else A.T) @ (B if ops[1] == "N" else B.T)
            errs = ["%.0e" % (ab.gemm(1, ops[0], A, ops[1], B) - ref).abs().max().item() for _ in range(4)]
            print("tile", tile, (m, n, k), ops, errs, flush=True)
