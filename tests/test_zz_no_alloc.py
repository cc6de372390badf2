"""accblas_dot_static: the dot product for hosts that cannot let the library allocate device memory.

Who needs it: a runtime that budgets all of HBM through its own memory manager (Accelerate's llvm-ptx backend tracks
every device array and evicts by that table, so memory a foreign call takes on the side is memory the manager still
counts as free), a process whose device memory has all been given to an arena / caching allocator (cudaMalloc just
fails there), and latency-sensitive callers for whom the device-wide synchronisation of cudaMalloc / cudaFree is the
problem.  accblas_dot keeps a small cudaMalloc'd slot per (thread, device, stream) or borrows from the library's pool;
accblas_dot_static keeps its per-CTA partial sums in a fixed 32 KiB array inside the library image instead.

What is checked here, through the raw C ABI (ctypes), on the GPU:
  * two calls in a row, on DIFFERENT inputs, each against accblas_dot on the same inputs (the tolerance is the one
    tests/test_parity_gpu.py::test_dot uses) and against the wide oracle -- the second call finds the first call's
    partial sums still in the shared array and must not be affected by them;
  * a third call that repeats the first call's inputs returns the first call's bits (fixed summation order);
  * the stated limits are refused with ACCBLAS_ERR_WORKSPACE and a sentence (CPU: the argument checks come before
    any CUDA call).

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure in this add-on must not hide
the parity tests.  It passes alone in a fresh process (then its first call does the one-time address look-up) and
after the rest of the suite.
"""
import ctypes

import numpy as np
import pytest

from oracle import oracle as O
from _util import TOL_UNITS

DTYPES = [O.F32, O.F64, O.C32, O.C64]
ids = lambda d: O.NAMES[d]  # noqa: E731
ERR_WORKSPACE = 7
MAX_N = 1 << 32        # ACCBLAS_DOT_STATIC_MAX_N (include/accblas_b200.h)


def test_limits_are_refused_with_a_sentence():
    import accblas_b200 as ab
    L = ab.lib()
    dummy = ctypes.c_void_p(256)                       # never dereferenced: the size check comes first
    st = L.accblas_dot_static(None, O.F64, 0, MAX_N + 1, dummy, dummy, dummy)
    assert st == ERR_WORKSPACE
    msg = L.accblas_last_cuda_error()
    msg = msg.decode() if isinstance(msg, bytes) else msg
    assert "accblas_dot_static" in msg and "ACCBLAS_DOT_STATIC_MAX_N" in msg
    assert L.accblas_dot_static(None, 9, 0, 4, dummy, dummy, dummy) == 1      # ACCBLAS_ERR_DTYPE
    assert L.accblas_dot_static(None, O.F64, 0, -1, dummy, dummy, dummy) == 3  # ACCBLAS_ERR_SHAPE
    assert L.accblas_dot_static(None, O.F64, 0, 4, dummy, dummy, None) == 4    # ACCBLAS_ERR_NULL


@pytest.mark.gpu
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_dot_static_matches_dot_on_two_calls_in_a_row(dt, cuda):
    import torch
    import accblas_b200 as ab
    L = ab.lib()
    rng = np.random.default_rng(9100 + dt)
    p = lambda t: ctypes.c_void_p(t.data_ptr())                                     # noqa: E731
    stream = ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream)
    complex_dt = dt in (O.C32, O.C64)

    def run(entry, conj, n, dx, dy):
        out = torch.full((1,), 7.0, dtype=dx.dtype, device=cuda)                   # a wrong value to overwrite
        st = entry(stream, dt, int(conj), n, p(dx), p(dy), p(out))
        assert st == 0, (st, L.accblas_last_cuda_error())
        return out

    # pairs of (first call, second call): big then small (the small launch uses fewer slots than the array still
    # holds from the big one), small then big, unaligned views, the empty dot
    for n1, n2 in [(1 << 22, 1000), (33, (1 << 21) + 3), (5_000_001, 0), (2049, 65537)]:
        for conj in ([False, True] if complex_dt else [False]):
            calls = []
            for n in (n1, n2):
                x, y = O.uniform(rng, (n + 1,), dt), O.uniform(rng, (n + 1,), dt)
                off = 1 if n == 2049 else 0                                        # 2049: operands off the 16-byte grid
                dx, dy = torch.from_numpy(x).to(cuda)[off:off + n], torch.from_numpy(y).to(cuda)[off:off + n]
                calls.append((n, x[off:off + n], y[off:off + n], dx, dy))
            # the two calls under test, back to back on one stream, nothing in between
            got = [run(L.accblas_dot_static, conj, n, dx, dy) for n, _, _, dx, dy in calls]
            again = run(L.accblas_dot_static, conj, calls[0][0], calls[0][3], calls[0][4])
            want = [run(L.accblas_dot, conj, n, dx, dy) for n, _, _, dx, dy in calls]
            torch.cuda.synchronize()
            for (n, x, y, _, _), g, w in zip(calls, got, want):
                g, w = complex(g.cpu().numpy()[0]), complex(w.cpu().numpy()[0])
                bound = (n + 2) * O.EPS[dt] * max(np.linalg.norm(x) * np.linalg.norm(y), 1e-300) * TOL_UNITS[dt]
                assert abs(g - w) <= bound, (O.NAMES[dt], n, conj, g, w)
                ref = complex(O.dot(x, y, conj=conj, wide=True))
                assert abs(g - ref) <= bound, (O.NAMES[dt], n, conj, g, ref)
            assert torch.equal(again, got[0]), (O.NAMES[dt], n1, conj)
