"""tools/exp/strip_ab.py <tree> [knob=value ...] -- the strip products of the host wavefront (dgemm 8192^3, 32 panels), device-resident,
timed one at a time and as the concurrent row + column pair, with the library of <tree>."""
import ctypes
import json
import os
import sys

tree = os.path.abspath(sys.argv[1])
sys.path.insert(0, os.path.join(tree, "accelerate-blas_b200", "python"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import accblas_b200 as ab  # noqa: E402

for kv in sys.argv[2:]:
    k, v = kv.split("=")
    ab.set_tuning(int(k), int(v))
n = 8192
L = ab.lib()
vp = ctypes.c_void_p
A = torch.rand((n, n), dtype=torch.float64, device="cuda")
B = torch.rand((n, n), dtype=torch.float64, device="cuda")
C = torch.empty((n, n), dtype=torch.float64, device="cuda")
one = np.array([1.0])
al = one.ctypes.data_as(vp)


def prod(stream, r0, h, c0, w):
    st = L.accblas_gemm_ld(vp(stream.cuda_stream), 1, 0, 0, h, n, n, w, al, vp(A.data_ptr() + 8 * r0 * n), n,
                           vp(B.data_ptr() + 8 * c0), n, vp(C.data_ptr() + 8 * (r0 * n + c0)), n)
    assert st == 0, st


s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
out = {"tree": tree, "knobs": sys.argv[2:]}


def timed(name, fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_event(e0); s2.wait_event(e0)
    for _ in range(iters):
        fn()
    ea, eb = torch.cuda.Event(), torch.cuda.Event()
    ea.record(s1); eb.record(s2)
    torch.cuda.current_stream().wait_event(ea); torch.cuda.current_stream().wait_event(eb)
    e1.record()
    torch.cuda.synchronize()
    out[name] = round(e0.elapsed_time(e1) / iters * 1e3, 1)      # us per iteration


for t in (0, 4, 12, 20, 30):
    r0 = c0 = t * 256
    if t:
        timed(f"row{t}", lambda: prod(s1, r0, 256, 0, t * 256))
    timed(f"col{t}", lambda: prod(s1, 0, (t + 1) * 256, c0, 256))
    if t:
        timed(f"pair{t}", lambda: (prod(s1, r0, 256, 0, t * 256), prod(s2, 0, (t + 1) * 256, c0, 256)))
# steps 28..31 back to back on two streams, as the wavefront issues them
def tail():
    for t in range(28, 32):
        prod(s1, t * 256, 256, 0, t * 256)
        prod(s2, 0, (t + 1) * 256, t * 256, 256)
timed("steps28_31", tail, iters=5)
print(json.dumps(out), flush=True)
