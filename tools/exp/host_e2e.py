"""tools/exp/host_e2e.py -- accblas_gemm_host dgemm 8192^3 (pinned buffers) under the knobs that changed in round 2."""
import ctypes
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
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


def run(reps=6):
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t = time.perf_counter()
        st = L.accblas_gemm_host(None, 1, 0, 0, n, n, n, n, one.ctypes.data_as(vp), vp(Ah.data_ptr()), vp(Bh.data_ptr()), vp(Ch.data_ptr()))
        assert st == 0
        ts.append((time.perf_counter() - t) * 1e3)
    ts.sort()
    return round(ts[len(ts) // 2], 3), round(ts[0], 3)


row = {"host_e2e": "dgemm_NN_8192", "pool": "default" if os.environ.get("ACCBLAS_DEFAULT_POOL") == "1" else "private"}
run(2)
row["auto"] = run()
for name, knob, val in (("onebox0", ab.TUNE_MN_ONEBOX, 0), ("release_fenced", ab.TUNE_RELEASE, 0), ("release_delayed", ab.TUNE_RELEASE, 2),
                        ("panels16", ab.TUNE_HOST_PANELS, 16), ("panels24", ab.TUNE_HOST_PANELS, 24), ("panels48", ab.TUNE_HOST_PANELS, 48)):
    ab.set_tuning(knob, val)
    row[name] = run()
    ab.set_tuning(knob, -1)
# raw link rates for the bound
torch.cuda.synchronize()
d = torch.empty((n, n), dtype=torch.float64, device="cuda")
for nm, fn in (("h2d", lambda: d.copy_(Ah, non_blocking=True)), ("d2h", lambda: Ch.copy_(d, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    t = time.perf_counter(); fn(); torch.cuda.synchronize()
    row[nm + "_gbs"] = round(8.0 * n * n / (time.perf_counter() - t) * 1e-9, 1)
print(json.dumps(row), flush=True)
