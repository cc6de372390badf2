"""GPU: round-3 fixes to existing entry points (the advisor's findings on round 2), each through the C ABI.

  * accblas_dot on a stream that is being captured into a CUDA graph takes the two-kernel path even when the stream
    already owns a single-launch slot from earlier eager calls (the slot's arrival counter must not be baked into a
    graph that may be replayed elsewhere); the captured graph replays correctly, also interleaved with eager dots;
  * the gemv T / H slab kernel with a thread-block cluster (a CTA announces itself before anybody touches its shared
    memory): shapes that select clusters of 2 and 4 against the oracle, repeated back to back;
  * D = beta * Cin with k == 0 through accblas_gemm_ex on the device (the validation order changed).
"""
import ctypes

import numpy as np
import pytest

from oracle import oracle as O
from _util import TOL_UNITS, gemv_inputs

pytestmark = pytest.mark.gpu
DTYPES = [O.F32, O.F64, O.C32, O.C64]
ids = lambda d: O.NAMES[d]  # noqa: E731


def test_dot_captured_on_a_stream_that_owns_a_slot(cuda):
    import torch
    import accblas_b200 as ab
    L = ab.lib()
    n = (1 << 20) + 7
    g = torch.Generator(device=cuda).manual_seed(5)
    x = torch.rand(n, dtype=torch.float64, device=cuda, generator=g) - 0.5
    y = torch.rand(n, dtype=torch.float64, device=cuda, generator=g) - 0.5
    p = lambda t: ctypes.c_void_p(t.data_ptr())                                     # noqa: E731
    s = torch.cuda.Stream(device=cuda)
    s.wait_stream(torch.cuda.current_stream(cuda))
    sp = ctypes.c_void_p(s.cuda_stream)
    eager = torch.zeros(1, dtype=torch.float64, device=cuda)
    out = torch.zeros(1, dtype=torch.float64, device=cuda)
    with torch.cuda.stream(s):
        L.accblas_set_tuning(21, 0)                                                 # one eager two-kernel dot first: the library's
        assert L.accblas_dot(sp, O.F64, 0, n, p(x), p(y), p(eager)) == 0            # memory pool exists before the capture starts
        L.accblas_set_tuning(21, -1)
        before = L.accblas_launch_count()
        assert L.accblas_dot(sp, O.F64, 0, n, p(x), p(y), p(eager)) == 0            # eager: the stream gets its slot
        assert L.accblas_launch_count() - before == 1                               # ... and the single-launch kernel
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            cs = ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream)
            before = L.accblas_launch_count()
            st = L.accblas_dot(cs, O.F64, 0, n, p(x), p(y), p(out))
            launches = L.accblas_launch_count() - before
        assert st == 0, L.accblas_last_cuda_error()
        assert launches == 2, "a captured dot must not use the stream's single-launch slot"
    s.synchronize()
    want = eager.clone()
    for rep in range(3):
        graph.replay()
        with torch.cuda.stream(s):                                                  # eager dots of the same stream beside it
            assert L.accblas_dot(sp, O.F64, 0, n, p(x), p(y), p(eager)) == 0
        torch.cuda.synchronize()
        assert torch.equal(out, want) and torch.equal(eager, want), rep
    x.mul_(2.0)                                                                     # the graph reads the live operands
    graph.replay()
    torch.cuda.synchronize()
    assert out.item() == 2.0 * want.item()


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemv_t_slab_clusters_back_to_back(dt, cuda):
    import accblas_b200 as ab
    rng = np.random.default_rng(9300 + dt)
    # 2000 x 3000 and 1000 x 2000 pick clusters of 2 / 4 (csrc/gemv.cu: slab_cluster); 129 x 4100: few rows per CTA
    for (m, n) in [(2000, 3000), (1000, 2000), (129, 4100), (513, 1030)]:
        for op in (["T", "H"] if dt in (O.C32, O.C64) else ["T"]):
            al, A, x = gemv_inputs(rng, dt, op, m, n)
            ref = O.gemv(al, op, A, x)
            first = None
            for rep in range(4):
                y = ab.gemv(al, op, A, x)
                assert O.gemm_error_units(y, ref, A, x, m, al) <= TOL_UNITS[dt], (O.NAMES[dt], m, n, op, rep)
                first = y if first is None else first
                assert np.array_equal(y, first)


def test_beta_times_cin_with_empty_inner_extent(cuda):
    import torch
    import accblas_b200 as ab
    L = ab.lib()
    m, n = 37, 53
    Cin = torch.arange(m * n, dtype=torch.float64, device=cuda).reshape(m, n)
    C = torch.full((m, n), -1.0, dtype=torch.float64, device=cuda)
    one, beta = np.array([1.0]), np.array([0.5])
    vp = ctypes.c_void_p
    st = L.accblas_gemm_ex(None, O.F64, 0, 0, m, 0, 0, n, one.ctypes.data_as(vp), None, 0, None, 0,
                           beta.ctypes.data_as(vp), vp(Cin.data_ptr()), 0, vp(C.data_ptr()), n)
    assert st == 0, L.accblas_last_cuda_error()
    torch.cuda.synchronize()
    assert torch.equal(C, 0.5 * Cin)
