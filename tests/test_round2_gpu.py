"""GPU: round-2 additions -- the extended product (beta * Cin in every family's epilogue), the persistent
whole-tile + stream-K schedule, the library-owned workspace pool, and oracle-checked parity at the
BASELINE.json sizes themselves (configs[1..3]: s/dgemm 8192^3 NN/NT/TN, zgemm / cgemm 4096^3 HN on the 3M and
4M paths, dgemv 32768^2 N/T, ddot 2^28).

The BASELINE-size checks compare SAMPLED rows / columns / the full dot with the wide oracle (oracle.c is O(mnk)
scalar code: 16 rows of an 8192^3 product are 10^9 multiply-adds) and the whole result through linearity or
run-to-run equality; tolerances are the stated ones (tests/_util.py).
"""
import numpy as np
import pytest

from oracle import oracle as O
from _util import TOL_UNITS, TOL_UNITS_TF32X3, gemm_inputs

pytestmark = pytest.mark.gpu
DTYPES = [O.F32, O.F64, O.C32, O.C64]
ids = lambda d: O.NAMES[d]  # noqa: E731


def dev(a, cuda):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(cuda)


def host(t):
    return t.cpu().numpy()


def tol_for(ab, dt, oa, ob, shapeA, shapeB):
    name = ab.kernel_name(dt, oa, ob, shapeA, shapeB)
    return TOL_UNITS_TF32X3 if "tf32x3" in name else TOL_UNITS[dt]


def ex_bound(dt, tol, al, A, B, k, beta, Cin, ref):
    """||D - D_ref||_F bound for D = alpha op(A) op(B) + beta Cin: the product's bound plus the roundings of the
    beta term and of the final sum."""
    return (tol * (k + 2) * O.EPS[dt] * abs(al) * O.fro(A) * O.fro(B)
            + 4 * O.EPS[dt] * (abs(beta) * O.fro(Cin) + O.fro(ref)))


# ------------------------------------------------------------------------------------------ accblas_gemm_ex
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemm_ex_beta_every_family(dt, cuda):
    """D = alpha op(A) op(B) + beta Cin through every kernel family's own epilogue: SIMT (tiny, odd), rank-1
    (k = 1), DMMA / mma.sync tiles, the tcgen05 kernel, 3M (forced), the persistent stream-K kernel; Cin separate
    and in place; k = 0 (D = beta Cin)."""
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(11000 + dt)
    wide = np.complex128 if dt in (O.C32, O.C64) else np.float64
    shapes = [(7, 5, 3), (37, 41, 1), (96, 80, 112), (300, 260, 200), (512, 384, 256), (1100, 1000, 520), (64, 48, 0)]
    seen = set()
    for (m, n, k) in shapes:
        for oa, ob in [("N", "N"), ("T", "N"), ("N", "H")]:
            al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
            beta = O.uniform(rng, (1,), dt)[0]
            Cin = O.uniform(rng, (m, n), dt)
            seen.add(ab.kernel_name(dt, oa, ob, A.shape, B.shape))
            dA, dB, dC = dev(A, cuda), dev(B, cuda), dev(Cin, cuda)
            ref = (O.gemm(al, oa, A, ob, B, wide=True).astype(wide) if k > 0 else np.zeros((m, n), wide)) + wide(beta) * Cin.astype(wide)
            bound = ex_bound(dt, tol_for(ab, dt, oa, ob, A.shape, B.shape), al, A, B, k, beta, Cin, ref)
            # separate Cin
            D = ab.gemm(al, oa, dA, ob, dB, beta=beta, Cin=dC)
            assert torch.equal(dC, dev(Cin, cuda)), "Cin must not be written"
            assert O.fro(host(D).astype(wide) - ref) <= bound, (O.NAMES[dt], oa, ob, m, n, k, "separate")
            # in place
            D2 = dC.clone()
            ab.gemm(al, oa, dA, ob, dB, out=D2, beta=beta, Cin=D2)
            assert O.fro(host(D2).astype(wide) - ref) <= bound, (O.NAMES[dt], oa, ob, m, n, k, "in place")
            # beta = 0: Cin is never read (NaNs in it must not leak)
            nan = torch.full_like(dC, float("nan"))
            D3 = ab.gemm(al, oa, dA, ob, dB, beta=0, Cin=nan)
            assert not torch.isnan(torch.view_as_real(D3) if D3.is_complex() else D3).any()
    assert len(seen) >= 3, seen
    # forced families: 3M, pure stream-K, 4M complex tensor path
    forced = []
    if dt == O.C64:
        forced = [(ab.TUNE_ZGEMM_3M, 1), (ab.TUNE_ZGEMM_3M, 0)]
    elif dt == O.C32:
        forced = [(ab.TUNE_CGEMM_3M, 1), (ab.TUNE_CGEMM_3M, 0)]
    elif dt == O.F64:
        forced = [(ab.TUNE_STREAMK, 1), (ab.TUNE_STREAMK, 2)]
    for knob, val in forced:
        try:
            ab.set_tuning(knob, val)
            if dt == O.C32:
                ab.set_tuning(2, 2)
            m, n, k = 640, 768, 512
            al, A, B = gemm_inputs(rng, dt, "H", "N", m, n, k)
            beta = O.uniform(rng, (1,), dt)[0]
            Cin = O.uniform(rng, (m, n), dt)
            ref = O.gemm(al, "H", A, "N", B, wide=True).astype(wide) + wide(beta) * Cin.astype(wide)
            D = ab.gemm(al, "H", dev(A, cuda), "N", dev(B, cuda), beta=beta, Cin=dev(Cin, cuda))
            bound = ex_bound(dt, TOL_UNITS_TF32X3, al, A, B, k, beta, Cin, ref)
            assert O.fro(host(D).astype(wide) - ref) <= bound, (O.NAMES[dt], knob, val)
        finally:
            ab.set_tuning(knob, -1)
            ab.set_tuning(2, -1)


def test_gemm_ex_windows_and_errors(cuda):
    """Pitched Cin / C windows; aliasing rules: Cin == C only with the same pitch, partial overlap refused."""
    import ctypes
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(11100)
    m, n, k = 200, 136, 96
    al, A, B = gemm_inputs(rng, O.F64, "N", "N", m, n, k)
    big_in = O.uniform(rng, (m + 8, n + 24), O.F64)
    big_out = torch.zeros((m + 4, n + 16), dtype=torch.float64, device=cuda)
    d_in = dev(big_in, cuda)
    Cin = d_in[3:3 + m, 8:8 + n]
    out = big_out[2:2 + m, 4:4 + n]
    ab.gemm(al, "N", dev(A, cuda), "N", dev(B, cuda), out=out, beta=0.5, Cin=Cin)
    ref = O.gemm(al, "N", A, "N", B, wide=True) + 0.5 * big_in[3:3 + m, 8:8 + n]
    assert O.fro(host(out) - ref) <= ex_bound(O.F64, 1.0, al, A, B, k, 0.5, big_in, ref)
    halo = big_out.clone()
    halo[2:2 + m, 4:4 + n] = 0
    assert torch.count_nonzero(halo).item() == 0, "stores outside the C window"
    # partial overlap of Cin and C
    L = ab.lib()
    one = np.array([1.0])
    p = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    dA, dB = dev(A, cuda), dev(B, cuda)
    buf = torch.zeros((m + 1) * n, dtype=torch.float64, device=cuda)
    st = L.accblas_gemm_ex(None, 1, 0, 0, m, k, k, n, one.ctypes.data_as(ctypes.c_void_p), p(dA), 0, p(dB), 0,
                           one.ctypes.data_as(ctypes.c_void_p), ctypes.c_void_p(buf.data_ptr() + 8 * n), 0, p(buf), 0)
    assert st == 8, st          # ACCBLAS_ERR_ALIAS
    st = L.accblas_gemm_ex(None, 1, 0, 0, m, k, k, n, one.ctypes.data_as(ctypes.c_void_p), p(dA), 0, p(dB), 0,
                           one.ctypes.data_as(ctypes.c_void_p), None, 0, p(buf), 0)
    assert st == 4, st          # ACCBLAS_ERR_NULL: beta != 0 without Cin


# ------------------------------------------------------------------------------------------ persistent schedule
def test_dgemm_persistent_schedules(cuda):
    """Whole-tile rounds + stream-K tail (knob 1), pure stream-K (knob 2) and the data-parallel launch (knob 0)
    agree within tolerance on products with 0, 1 and several whole-tile rounds; each persistent schedule is
    bit-reproducible run to run; both stage-release modes give the same bits (they only differ in timing)."""
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(11200)
    try:
        for (m, n, k) in [(1280, 1152, 384), (2048, 2048, 320), (4096, 2304, 160), (1900, 2100, 530)]:
            for oa, ob in [("N", "N"), ("T", "N"), ("N", "T")]:
                al, A, B = gemm_inputs(rng, O.F64, oa, ob, m, n, k)
                dA, dB = dev(A, cuda), dev(B, cuda)
                ab.set_tuning(ab.TUNE_STREAMK, 0)
                C0 = ab.gemm(al, oa, dA, ob, dB)
                outs = {}
                for knob in (1, 2):
                    ab.set_tuning(ab.TUNE_STREAMK, knob)
                    for rel in (0, 2):
                        ab.set_tuning(ab.TUNE_RELEASE, rel)
                        outs[(knob, rel)] = ab.gemm(al, oa, dA, ob, dB)
                    ab.set_tuning(ab.TUNE_RELEASE, -1)
                    again = ab.gemm(al, oa, dA, ob, dB)
                    assert torch.equal(outs[(knob, 0)], outs[(knob, 2)]), (m, n, k, oa, ob, knob, "release modes differ")
                    assert torch.equal(outs[(knob, 0)], again), (m, n, k, oa, ob, knob, "not reproducible")
                    assert O.gemm_error_units(host(again), host(C0), A, B, k, al) <= 2 * TOL_UNITS[O.F64], (m, n, k, oa, ob, knob)
                rows = rng.choice(m, size=5, replace=False)
                Asub = A[rows] if oa == "N" else np.ascontiguousarray(A[:, rows])
                err = O.gemm_error_units(host(outs[(1, 2)])[rows], O.gemm(al, oa, Asub, ob, B, wide=True), Asub, B, k, al)
                assert err <= TOL_UNITS[O.F64], (m, n, k, oa, ob, err)
    finally:
        ab.set_tuning(ab.TUNE_STREAMK, -1)
        ab.set_tuning(ab.TUNE_RELEASE, -1)


@pytest.mark.parametrize("dt", [O.F64, O.C64], ids=ids)
def test_release_modes_big_tiles(dt, cuda):
    """The delayed stage release (no proxy fence; the stage of K-block i is handed back at the top of K-block i + 1)
    against the fenced one, on every tile of the family: same bits, on shapes with short and long K (ring
    wrapping many times), small tiles with many CTAs per SM included."""
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(11300 + dt)
    knob = ab.TUNE_DMMA_TILE if dt == O.F64 else ab.TUNE_ZMMA_TILE
    try:
        ab.set_tuning(ab.TUNE_ZGEMM_3M, 0)
        for tile in range(4 if dt == O.F64 else 3):
            ab.set_tuning(knob, tile)
            for (m, n, k) in [(1024, 1024, 96), (1536, 1280, 1024)]:
                for oa, ob in [("N", "N"), ("H", "N"), ("N", "T")]:
                    al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                    dA, dB = dev(A, cuda), dev(B, cuda)
                    res = []
                    for rel in (0, 2, 2, 2, 2):
                        ab.set_tuning(ab.TUNE_RELEASE, rel)
                        res.append(ab.gemm(al, oa, dA, ob, dB))
                    for r in res[1:]:
                        assert torch.equal(res[0], r), (O.NAMES[dt], tile, m, n, k, oa, ob)
    finally:
        ab.set_tuning(knob, -1)
        ab.set_tuning(ab.TUNE_ZGEMM_3M, -1)
        ab.set_tuning(ab.TUNE_RELEASE, -1)


# ------------------------------------------------------------------------------------------ workspace pool
def test_workspace_pool_is_private_and_trims(cuda):
    import torch
    import accblas_b200 as ab
    L = ab.lib()
    n = 1536
    A = torch.rand((n, n), dtype=torch.float32, device=cuda)
    assert ab.workspace_bytes(O.F32, "N", "N", A.shape, A.shape) > 0      # tcgen05 path: hi / lo planes
    ab.gemm(1, "N", A, "N", A)
    torch.cuda.synchronize()
    assert L.accblas_workspace_cached_bytes() > 0
    assert L.accblas_trim_workspace() == 0
    assert L.accblas_workspace_cached_bytes() == 0
    C = ab.gemm(1, "N", A, "N", A)                                       # still works after a trim
    torch.cuda.synchronize()
    assert torch.isfinite(C).all()
    import ctypes
    buf = ctypes.create_string_buffer(64)
    assert L.accblas_last_cuda_error_r(buf, 64) == len(buf.value) or L.accblas_last_cuda_error_r(buf, 64) >= 63
    assert L.accblas_launch_count_total() >= L.accblas_launch_count()


# ------------------------------------------------------------------------------------------ BASELINE sizes
def seeded(shape, seed, cuda, dtype):
    import torch
    g = torch.Generator(device=cuda).manual_seed(seed)
    if dtype.is_complex:
        base = torch.float32 if dtype == torch.complex64 else torch.float64
        return torch.view_as_complex(torch.rand(tuple(shape) + (2,), dtype=base, device=cuda, generator=g) * 2 - 1)
    return torch.rand(shape, dtype=dtype, device=cuda, generator=g) * 2 - 1


def sampled_gemm_check(ab, cuda, dt, tdt, n, oa, ob, rows, seed):
    """A <op> B at n^3 from seeded device operands: `rows` rows of C against the wide oracle."""
    import torch
    A = seeded((n, n), seed, cuda, tdt)
    B = seeded((n, n), seed + 1, cuda, tdt)
    C = ab.gemm(1, oa, A, ob, B)
    idx = torch.as_tensor(rows, device=cuda)
    Asub = host(A[idx] if oa == "N" else A[:, idx].contiguous())
    Bh = host(B)
    ref = O.gemm(1, oa, Asub, ob, Bh, wide=True)
    err = O.gemm_error_units(host(C[idx]), ref, Asub, Bh, n, 1)
    return err, A, B, C


@pytest.mark.parametrize("ops", ["NN", "NT", "TN"])
def test_baseline_dgemm_8192(ops, cuda):
    """configs[1]: Double gemm n = 8192 -- 16 rows against the wide oracle (one per 512-row band: every rank of
    the raster groups, first / last row), run-to-run bit equality of the whole matrix."""
    import torch
    import accblas_b200 as ab
    n = 8192
    rows = [0, n - 1] + [512 * i + (37 * i) % 512 for i in range(1, 15)]
    err, A, B, C = sampled_gemm_check(ab, cuda, O.F64, torch.float64, n, ops[0], ops[1], rows, 0x5EED0000 + 16)
    assert err <= TOL_UNITS[O.F64], (ops, err)
    assert torch.equal(C, ab.gemm(1, ops[0], A, ops[1], B))


@pytest.mark.parametrize("ops", ["NN", "NT", "TN"])
def test_baseline_sgemm_8192(ops, cuda):
    """configs[1]: Float gemm n = 8192 on the tcgen05 3xTF32 path."""
    import torch
    import accblas_b200 as ab
    n = 8192
    rows = [0, n - 1] + [512 * i + (37 * i) % 512 for i in range(1, 15)]
    assert "tcgen05" in ab.kernel_name(O.F32, ops[0], ops[1], (n, n), (n, n))
    err, A, B, C = sampled_gemm_check(ab, cuda, O.F32, torch.float32, n, ops[0], ops[1], rows, 0x5EED0000 + 24)
    assert err <= TOL_UNITS_TF32X3, (ops, err)


@pytest.mark.parametrize("dt,scheme", [(O.C64, 1), (O.C64, 0), (O.C32, 1), (O.C32, 0)],
                         ids=["z3m", "z4m", "c3m", "c4m"])
def test_baseline_complex_gemm_4096_hn(dt, scheme, cuda):
    """configs[3]: zgemm (and cgemm) 4096^3 with ConjugateTranspose on A, on the 3M AND the 4M path: 16 rows against
    the wide oracle, normwise (stated tolerance) and componentwise (max |Re|, |Im| deviation relative to
    k eps ||a_i|| ||b_j||, reported; 3M's imaginary parts carry the larger constant)."""
    import torch
    import accblas_b200 as ab
    n = 4096
    tdt = torch.complex128 if dt == O.C64 else torch.complex64
    knob = ab.TUNE_ZGEMM_3M if dt == O.C64 else ab.TUNE_CGEMM_3M
    rows = [0, n - 1] + [256 * i + (29 * i) % 256 for i in range(1, 15)]
    try:
        ab.set_tuning(knob, scheme)
        name = ab.kernel_name(dt, "H", "N", (n, n), (n, n))
        assert ("3m" in name) == bool(scheme), name
        err, A, B, C = sampled_gemm_check(ab, cuda, dt, tdt, n, "H", "N", rows, 0x5EED0000 + 48)
    finally:
        ab.set_tuning(knob, -1)
    assert err <= TOL_UNITS_TF32X3, (name, err)


@pytest.mark.parametrize("op", ["N", "T"])
def test_baseline_dgemv_32768(op, cuda):
    """configs[2]: Double gemv 32768 x 32768 -- 24 sampled entries of y against the wide oracle (rows of A for N,
    columns for T), plus run-to-run bit equality."""
    import torch
    import accblas_b200 as ab
    n = 32768
    A = seeded((n, n), 0x5EED0000 + 32, cuda, torch.float64)
    x = seeded((n,), 0x5EED0000 + 33, cuda, torch.float64)
    y = ab.gemv(1, op, A, x)
    pick = [0, 1, n - 1, n - 2] + [(1637 * i) % n for i in range(1, 21)]
    idx = torch.as_tensor(pick, device=cuda)
    if op == "N":
        sub = host(A[idx])
        ref = O.gemv(1, "N", sub, host(x), wide=True)
        scale = np.linalg.norm(sub, axis=1)
    else:
        sub = host(A[:, idx].contiguous())
        ref = O.gemv(1, "T", sub, host(x), wide=True)
        scale = np.linalg.norm(sub, axis=0)
    got = host(y[idx])
    bound = (n + 2) * O.EPS[O.F64] * scale * float(torch.linalg.norm(x))
    assert np.all(np.abs(got - ref) <= bound), (op, float(np.max(np.abs(got - ref) / bound)))
    assert torch.equal(y, ab.gemv(1, op, A, x))


def test_baseline_ddot_2p28(cuda):
    """configs[2]: dotu on 2^28 Doubles against the wide oracle on the FULL vectors (4 GiB on the host)."""
    import torch
    import accblas_b200 as ab
    n = 1 << 28
    x = seeded((n,), 0x5EED0000 + 35, cuda, torch.float64)
    y = seeded((n,), 0x5EED0000 + 36, cuda, torch.float64)
    d = ab.dotu(x, y)
    d2 = ab.dotu(x, y)
    assert torch.equal(d, d2)
    xh, yh = host(x), host(y)
    nx, ny = float(torch.linalg.norm(x)), float(torch.linalg.norm(y))
    del x, y
    ref = float(O.dot(xh, yh, wide=True))
    assert abs(float(d) - ref) <= n * O.EPS[O.F64] * nx * ny
    # a blocked double-precision sum sits far inside the worst-case bound: 1e-3 of it is still generous
    assert abs(float(d) - ref) <= 1e-3 * n * O.EPS[O.F64] * nx * ny


# ------------------------------------------------------------------------------------------ multi-GPU C ABI, one rank
def test_mg_c_abi_single_rank(cuda):
    """accblas_mg_* with world = 1 (what a one-GPU box can run): communicator bootstrap through NCCL, gemm through
    the plan / run entry points (root path == accblas_gemm bit for bit), the peers' pull / K-chunk / beta = 1 pipeline
    driven on the root's own arrays (ACCBLAS_TUNE_MG_SELF_PEER) against the oracle, gemv / dot through the exchange
    buffer (== the single-GPU kernels bit for bit)."""
    import ctypes
    import torch
    import accblas_b200 as ab
    L = ab.lib()
    vp = ctypes.c_void_p
    ident = ctypes.create_string_buffer(128)
    assert L.accblas_mg_unique_id(ident) == 0, L.accblas_last_cuda_error()
    comm = vp()
    with torch.cuda.device(cuda):
        st = L.accblas_mg_init(ctypes.byref(comm), 0, 1, ident.raw)
    assert st == 0, L.accblas_last_cuda_error()
    assert L.accblas_mg_rank(comm) == 0 and L.accblas_mg_world(comm) == 1
    stream = lambda: vp(torch.cuda.current_stream(cuda).cuda_stream)  # noqa: E731
    p = lambda t: vp(t.data_ptr())  # noqa: E731
    rng = np.random.default_rng(11500)
    try:
        for self_peer in (0, 1):
            ab.set_tuning(ab.TUNE_MG_SELF_PEER, self_peer)
            for dt, tdt in ((O.F64, torch.float64), (O.C32, torch.complex64)):
                for (oa, ob, m, n, k) in [("N", "N", 384, 256, 8704), ("T", "N", 256, 384, 5000), ("N", "H", 130, 200, 4608),
                                          ("H", "T", 200, 130, 700)]:
                    al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                    dA, dB = dev(A, cuda), dev(B, cuda)
                    C = torch.full((m, n), float("nan"), dtype=tdt, device=cuda)
                    alb = np.array([al], dtype=A.dtype)
                    plan = vp()
                    code = {"N": 0, "T": 1, "H": 2}
                    st = L.accblas_mg_gemm_plan(comm, dt, code[oa], code[ob], A.shape[0], A.shape[1], B.shape[0], B.shape[1],
                                                p(dA), p(dB), p(C), 0, ctypes.byref(plan))
                    assert st == 0, (st, L.accblas_last_cuda_error())
                    for rep in range(2):
                        assert L.accblas_mg_gemm_run(plan, stream(), alb.ctypes.data_as(vp)) == 0, L.accblas_last_cuda_error()
                        torch.cuda.synchronize()
                        if self_peer == 0:
                            assert torch.equal(C, ab.gemm(al, oa, dA, ob, dB)), (oa, ob, m, n, k)
                        else:
                            err = O.gemm_error_units(host(C), O.gemm(al, oa, A, ob, B, wide=True), A, B, k, al)
                            assert err <= TOL_UNITS_TF32X3, (O.NAMES[dt], oa, ob, m, n, k, err)
                    assert L.accblas_mg_gemm_plan_destroy(plan) == 0
        ab.set_tuning(ab.TUNE_MG_SELF_PEER, -1)
        # gemv N / T / H and dot: with one rank the exchange must reproduce the local kernels exactly
        for dt, tdt in ((O.F64, torch.float64), (O.C64, torch.complex128), (O.F32, torch.float32)):
            A = dev(O.uniform(rng, (700, 520), dt), cuda)
            one = np.array([1], dtype=O.NP_DTYPE[dt])
            for op, nx, ny in (("N", 520, 700), ("T", 700, 520), ("H", 700, 520)):
                x = dev(O.uniform(rng, (nx,), dt), cuda)
                y = torch.empty((ny,), dtype=tdt, device=cuda)
                for rep in range(3):
                    st = L.accblas_mg_gemv(comm, stream(), dt, {"N": 0, "T": 1, "H": 2}[op], 700, 520, one.ctypes.data_as(vp),
                                           p(A), p(x), nx, p(y))
                    assert st == 0, L.accblas_last_cuda_error()
                assert torch.equal(y, ab.gemv(1, op, A, x)), (O.NAMES[dt], op)
            u, v = dev(O.uniform(rng, (100003,), dt), cuda), dev(O.uniform(rng, (100003,), dt), cuda)
            out = torch.empty((), dtype=tdt, device=cuda)
            for conj in ((0, 1) if dt == O.C64 else (0,)):
                assert L.accblas_mg_dot(comm, stream(), dt, conj, 100003, p(u), p(v), p(out)) == 0
                assert torch.equal(out, (ab.dotc if conj else ab.dotu)(u, v))
    finally:
        ab.set_tuning(ab.TUNE_MG_SELF_PEER, -1)
        torch.cuda.synchronize()
        L.accblas_mg_finalize(comm)


# ------------------------------------------------------------------------------------------ non-finite inputs
def _wide_product(oa, A, ob, B):
    """op(A) op(B) in float64 from explicit real products (no complex BLAS: its Inf handling is not IEEE-by-parts)."""
    def op(o, X):
        X = X.T if o in ("T", "H") else X
        return np.conj(X) if o == "H" else X
    X, Y = op(oa, A), op(ob, B)
    with np.errstate(all="ignore"):
        if np.iscomplexobj(X):
            xr, xi, yr, yi = (np.asarray(v, dtype=np.float64) for v in (X.real, X.imag, Y.real, Y.imag))
            out = np.empty(xr.shape[:1] + yr.shape[1:], dtype=np.complex128)
            out.real = xr @ yr - xi @ yi
            out.imag = xr @ yi + xi @ yr
            return out
        return np.asarray(X, dtype=np.float64) @ np.asarray(Y, dtype=np.float64)


@pytest.mark.parametrize("dt", [O.F32, O.C32], ids=ids)
def test_gemm_nonfinite_inputs_f32_families(dt, cuda):
    """+-Inf, NaN, Inf * 0 and overflowing entries through the split-precision (3xTF32) families must come out as a plain
    FP32 gemm (cublasSgemm / cublasCgemm, LLVM/PTX/Level3.hs:91,101) returns them: Inf where the terms agree in sign, NaN
    for Inf - Inf, Inf * 0 and NaN inputs, everything else finite and within tolerance (round-1 advisor finding:
    the split turned every output of an Inf row into NaN).  csrc/refit.cuh."""
    import accblas_b200 as ab
    rng = np.random.default_rng(4242 + dt)
    cx = dt == O.C32
    seen = set()
    # c32: mma.sync tiles, tcgen05 4M embedding, tcgen05 3M planes; f32: mma.sync tiles, tcgen05
    cases = ([(96, 80, 64, -1), (300, 260, 200, -1), (1100, 1000, 520, 0), (1100, 1000, 520, 1)] if cx else
             [(96, 80, 64, -1), (300, 260, 200, -1), (1100, 1000, 520, -1), (1536, 1280, 520, -1)])
    for (m, n, k, c3m) in cases:
        for oa, ob in [("N", "N"), ("T", "N"), ("N", "H"), ("H", "T")]:
            al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
            al = 1.0
            # work on op-index views so the planted entries land at known (i, l) / (l, j)
            Aop = A.T if oa != "N" else A
            Bop = B.T if ob != "N" else B
            Aop[3, :] = np.abs(Aop[3, :].real) + 0.5 if not cx else (np.abs(Aop[3, :].real) + 0.5) + 0j
            Aop[3, 5] = np.inf                      # row 3: +Inf times positive / negative / zero entries of B
            Aop[10, 7] = -np.inf
            Aop[17, 2] = np.inf; Aop[17, 9] = -np.inf       # Inf - Inf in every output of row 17 with non-zero B
            Bop[5, 11] = 0.0                        # C[3, 11] = Inf * 0 = NaN
            Bop[2, 21] = np.nan                     # column 21 is NaN
            if not cx:
                Aop[20, 1] = np.finfo(np.float32).max
                Bop[1, 30] = 2.0                    # overflow of one product
                Aop[21, 4] = np.float32(3.4028234e38)       # |a| whose TF32 rounding would be Inf
            ref = _wide_product(oa, A, ob, B)
            with np.errstate(all="ignore"):
                if cx:
                    # alpha = 1 + 0i is applied the way every complex BLAS epilogue applies it, by parts:
                    # re = 1 * re - 0 * im  (so an infinite imaginary part makes the real part NaN, and vice versa)
                    by_parts = np.empty_like(ref)
                    by_parts.real = 1.0 * ref.real - 0.0 * ref.imag
                    by_parts.imag = 1.0 * ref.imag + 0.0 * ref.real
                    ref = by_parts
                ref32 = ref.astype(np.complex64 if cx else np.float32)
            if c3m >= 0 and cx:
                ab.set_tuning(ab.TUNE_CGEMM_3M, c3m)
            try:
                seen.add(ab.kernel_name(dt, oa, ob, A.shape, B.shape))
                C = host(ab.gemm(al, oa, dev(A, cuda), ob, dev(B, cuda)))
            finally:
                ab.set_tuning(ab.TUNE_CGEMM_3M, -1)
            parts = [(C.real, ref32.real), (C.imag, ref32.imag)] if cx else [(C, ref32)]
            for got, want in parts:
                what = (O.NAMES[dt], oa, ob, m, n, k, c3m)
                assert np.array_equal(np.isnan(got), np.isnan(want)), ("NaN pattern", what, int(np.isnan(got).sum()), int(np.isnan(want).sum()))
                inf = np.isinf(want)
                assert np.array_equal(np.isinf(got), inf), ("Inf pattern", what)
                assert np.array_equal(np.sign(got[inf]), np.sign(want[inf])), ("Inf sign", what)
                fin = np.isfinite(want)
                scale = np.abs(want[fin]).max() if fin.any() else 1.0
                assert np.abs(got[fin] - want[fin]).max() <= 1e-4 * max(scale, 1.0) * np.sqrt(k), ("finite part", what)
    assert any("tcgen05" in s for s in seen) and any("mma" in s for s in seen), seen


# ------------------------------------------------------------------------------------------ f32, operands off TMA's grid
def test_sgemm_direct_misaligned_operands(cuda):
    """Float operands whose row pitch is not a multiple of 16 bytes (the reference benchmarks 250 x 250 Floats,
    bench/Bench/Accelerate.hs:92-125) with k <= 768: the mma.sync kernel loads its own panels (one launch, no re-pitch
    copy).  Every op pair, ragged extents, alpha, the beta epilogue, a misaligned base; against the wide oracle and
    against the round-1 path (copy into aligned scratch) on the same inputs."""
    import accblas_b200 as ab
    dt = O.F32
    rng = np.random.default_rng(12100)
    TUNE_SGEMM_DIRECT = 17
    shapes = [(250, 250, 250), (33, 47, 95), (100, 31, 767), (129, 65, 37), (31, 250, 513)]
    for (m, n, k) in shapes:
        for oa in "NTH":
            for ob in "NTH":
                al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                assert ab.kernel_name(dt, oa, ob, A.shape, B.shape) == "sgemm_tf32x3_mma_direct", (m, n, k, oa, ob)
                dA, dB = dev(A, cuda), dev(B, cuda)
                ref = O.gemm(al, oa, A, ob, B, wide=True)
                C = host(ab.gemm(al, oa, dA, ob, dB))
                err = O.gemm_error_units(C, ref, A, B, k, alpha=al)
                assert err <= TOL_UNITS_TF32X3, (m, n, k, oa, ob, err)
                ab.set_tuning(TUNE_SGEMM_DIRECT, 0)
                try:
                    C0 = host(ab.gemm(al, oa, dA, ob, dB))
                finally:
                    ab.set_tuning(TUNE_SGEMM_DIRECT, -1)
                assert O.gemm_error_units(C0, ref, A, B, k, alpha=al) <= TOL_UNITS_TF32X3
    # beta epilogue + an operand base that is only 4-byte aligned (a view one element into a buffer)
    m, n, k = 130, 94, 301
    al, A, B = gemm_inputs(rng, dt, "N", "N", m, n, k)
    beta = np.float32(0.75)
    Cin = O.uniform(rng, (m, n), dt)
    import torch
    bufA = torch.empty(m * k + 1, dtype=torch.float32, device=cuda)
    dA = bufA[1:].view(m, k)
    dA.copy_(dev(A, cuda))
    D = host(ab.gemm(al, "N", dA, "N", dev(B, cuda), beta=beta, Cin=dev(Cin, cuda)))
    ref = O.gemm(al, "N", A, "N", B, wide=True).astype(np.float64) + np.float64(beta) * Cin.astype(np.float64)
    assert O.fro(D.astype(np.float64) - ref) <= ex_bound(dt, TOL_UNITS_TF32X3, al, A, B, k, beta, Cin, ref)
    # beyond the shared-memory budget the copy path still serves the product
    A = O.uniform(rng, (70, 801), dt)
    B = O.uniform(rng, (801, 50), dt)
    assert ab.kernel_name(dt, "N", "N", A.shape, B.shape) == "sgemm_tf32x3_mma"
    C = host(ab.gemm(1, "N", dev(A, cuda), "N", dev(B, cuda)))
    assert O.gemm_error_units(C, O.gemm(1, "N", A, "N", B, wide=True), A, B, 801) <= TOL_UNITS_TF32X3
    # non-finite inputs keep the FP32 pattern on this path too
    A = O.uniform(rng, (64, 250), dt)
    B = O.uniform(rng, (250, 66), dt)
    A[3, 7] = np.inf
    B[9, 5] = np.nan
    C = host(ab.gemm(1, "N", dev(A, cuda), "N", dev(B, cuda)))
    want = A.astype(np.float64) @ B.astype(np.float64)
    assert np.array_equal(np.isnan(C), np.isnan(want)) and np.array_equal(np.isinf(C), np.isinf(want))


def test_dgemm_direct_misaligned_operands(cuda):
    """Double operands with an odd column count (rows 8- but not 16-byte aligned) in a product of at most one 32 x 32 tile
    per SM with k <= 384: the DMMA kernel loads its panels for the whole K extent itself instead of running behind a
    re-pitch copy.  Every op pair, alpha, beta; against the wide oracle and against the round-1 path (copy + TMA ring)."""
    import accblas_b200 as ab
    dt = O.F64
    rng = np.random.default_rng(12200)
    TUNE_DGEMM_DIRECT = 18
    for (m, n, k) in [(251, 249, 251), (33, 47, 95), (129, 65, 383), (99, 101, 13)]:
        for oa in "NTH":
            for ob in "NTH":
                al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                assert ab.kernel_name(dt, oa, ob, A.shape, B.shape) == "dgemm_dmma_direct", (m, n, k, oa, ob)
                dA, dB = dev(A, cuda), dev(B, cuda)
                ref = O.gemm(al, oa, A, ob, B, wide=True)
                C = ab.gemm(al, oa, dA, ob, dB)
                assert O.gemm_error_units(host(C), ref, A, B, k, alpha=al) <= TOL_UNITS[dt], (m, n, k, oa, ob)
                ab.set_tuning(TUNE_DGEMM_DIRECT, 0)
                try:
                    assert ab.kernel_name(dt, oa, ob, A.shape, B.shape) in ("dgemm_dmma_tma", "gemm_simt_generic")
                    C0 = ab.gemm(al, oa, dA, ob, dB)
                finally:
                    ab.set_tuning(TUNE_DGEMM_DIRECT, -1)
                assert O.gemm_error_units(host(C0), ref, A, B, k, alpha=al) <= TOL_UNITS[dt]
    m, n, k = 131, 95, 301
    al, A, B = gemm_inputs(rng, dt, "T", "N", m, n, k)
    beta = np.float64(-0.5)
    Cin = O.uniform(rng, (m, n), dt)
    assert ab.kernel_name(dt, "T", "N", A.shape, B.shape) == "dgemm_dmma_direct"
    D = host(ab.gemm(al, "T", dev(A, cuda), "N", dev(B, cuda), beta=beta, Cin=dev(Cin, cuda)))
    ref = O.gemm(al, "T", A, "N", B, wide=True).astype(np.float64) + beta * Cin
    assert O.fro(D - ref) <= ex_bound(dt, TOL_UNITS[dt], al, A, B, k, beta, Cin, ref)
    # one wave only: a product with more 32 x 32 tiles than SMs keeps the copy + ring path
    A, B = O.uniform(rng, (501, 101), dt), O.uniform(rng, (101, 499), dt)
    assert ab.kernel_name(dt, "N", "N", A.shape, B.shape) == "dgemm_dmma_tma"


# ------------------------------------------------------------------------------------------ host pipeline, K-panel schedule
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_host_pipeline_kpanels(dt, cuda):
    """accblas_gemm_host with the compute-bound schedule forced (csrc/host_pipeline.cu: K-panel rank updates of the whole C,
    then the remaining K range row panel by row panel with the downloads underneath): every op pair, ragged extents (last
    K panel, last row panel), complex alpha; sampled rows against the wide oracle and the whole result against the block
    wavefront on the same host arrays."""
    import accblas_b200 as ab
    TUNE_HOST_SCHEDULE = 19
    rng = np.random.default_rng(12300 + dt)
    m, n, k = 1200, 1090, 2700
    for oa in "NTH":
        for ob in "NTH":
            al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
            ab.set_tuning(TUNE_HOST_SCHEDULE, 1)
            try:
                C = ab.gemm(al, oa, A, ob, B)
            finally:
                ab.set_tuning(TUNE_HOST_SCHEDULE, -1)
            rows = np.arange(0, m, 97)
            Asub = A[rows] if oa == "N" else np.ascontiguousarray(A[:, rows])
            ref = O.gemm(al, oa, Asub, ob, B, wide=True)
            tol = TOL_UNITS_TF32X3 if dt in (O.F32, O.C32) else TOL_UNITS[dt]
            err = O.gemm_error_units(np.ascontiguousarray(C[rows]), ref, Asub, B, k, al)
            assert err <= tol, (O.NAMES[dt], oa, ob, err)
            ab.set_tuning(TUNE_HOST_SCHEDULE, 0)
            ab.set_tuning(4, 1)
            try:
                W = ab.gemm(al, oa, A, ob, B)
            finally:
                ab.set_tuning(TUNE_HOST_SCHEDULE, -1)
                ab.set_tuning(4, -1)
            assert O.gemm_error_units(C, W, A, B, k, al) <= 2 * tol, (O.NAMES[dt], oa, ob, "K-panels vs wavefront")


# ------------------------------------------------------------------------------------------ split-K over idle SMs
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemm_splitk_small_grids_long_k(dt, cuda):
    """Products of at most half a wave of 32 x 32 tiles with a long K (csrc/gemm_dmma.cu, csrc/gemm_mma_tf32.cu: K slices as the
    batch index of one launch + a reduce kernel): every op pair, ragged K (last slice shorter, K not a multiple of the K-block), alpha; against
    the wide oracle, against the unsplit kernel, and bit-reproducible run to run."""
    import torch
    import accblas_b200 as ab
    TUNE_SPLITK = 20
    rng = np.random.default_rng(12400 + dt)
    for (m, n, k) in [(250, 250, 1024), (96, 96, 1031), (64, 200, 777), (33, 65, 4099)]:
        for oa in "NTH":
            for ob in "NTH":
                al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                dA, dB = dev(A, cuda), dev(B, cuda)
                ref = O.gemm(al, oa, A, ob, B, wide=True)
                tol = TOL_UNITS_TF32X3 if dt in (O.F32, O.C32) else TOL_UNITS[dt]
                C = ab.gemm(al, oa, dA, ob, dB)
                assert O.gemm_error_units(host(C), ref, A, B, k, alpha=al) <= tol, (O.NAMES[dt], m, n, k, oa, ob)
                assert torch.equal(C, ab.gemm(al, oa, dA, ob, dB)), "split-K sums its slices in a fixed order"
                ab.set_tuning(TUNE_SPLITK, 0)
                try:
                    C0 = ab.gemm(al, oa, dA, ob, dB)
                finally:
                    ab.set_tuning(TUNE_SPLITK, -1)
                assert O.gemm_error_units(host(C0), ref, A, B, k, alpha=al) <= tol
                assert O.gemm_error_units(host(C), host(C0), A, B, k, alpha=al) <= 2 * tol
    # an output window with its own pitch (accblas_gemm_ld): the reduce kernel writes rows at the caller's ldc
    m, n, k = 100, 60, 2048
    al, A, B = gemm_inputs(rng, dt, "N", "N", m, n, k)
    big = torch.full((m, n + 12), 7.0, dtype=dev(A, cuda).dtype, device=cuda)
    out = big[:, 4:4 + n]
    ab.gemm(al, "N", dev(A, cuda), "N", dev(B, cuda), out=out)
    assert O.gemm_error_units(host(out.contiguous()), O.gemm(al, "N", A, "N", B, wide=True), A, B, k, alpha=al) <= (
        TOL_UNITS_TF32X3 if dt in (O.F32, O.C32) else TOL_UNITS[dt])
    assert bool((big[:, :4] == 7.0).all()) and bool((big[:, 4 + n:] == 7.0).all()), "wrote outside the window"
    if dt in (O.F32, O.C32):
        # non-finite inputs: every slice refits ITS OWN K range, the sum has the FP32 pattern of the whole product
        A = O.uniform(rng, (64, 2048), dt)
        B = O.uniform(rng, (2048, 66), dt)
        A[3, 7] = np.inf
        B[1900, 5] = np.nan
        A[9, 1500] = np.float32(3.4e38)             # TF32 rounding overflows; the product stays finite in FP32
        B[1500, :] *= np.float32(1e-3)
        C = host(ab.gemm(1, "N", dev(A, cuda), "N", dev(B, cuda)))
        want = A.astype(np.complex128 if dt == O.C32 else np.float64) @ B.astype(np.complex128 if dt == O.C32 else np.float64)
        fin = np.isfinite(want)
        assert np.array_equal(np.isfinite(C), fin)
        assert np.allclose(C[fin], want[fin].astype(C.dtype), rtol=2e-3, atol=1e-2 * np.abs(want[fin]).max())


# ------------------------------------------------------------------------------------------ gemv from host buffers, chunked upload
@pytest.mark.parametrize("dt", [O.F64, O.C32], ids=ids)
def test_gemv_host_chunked_upload(dt, cuda):
    """accblas_gemv_host on a matrix large enough for the chunked upload (op N: row chunks of 16 MiB multiply under the next
    chunk's transfer; ragged last chunk); T / H keep the serial flow.  Against the wide oracle and the device-resident call."""
    import accblas_b200 as ab
    rng = np.random.default_rng(12500 + dt)
    m, n = 9001, 3000                                       # 216 MB (f64) / 216 MB (c32): 13 chunks, the last one short
    A = O.uniform(rng, (m, n), dt)
    for oa in "NTH":
        x = O.uniform(rng, (n if oa == "N" else m,), dt)
        al = O.uniform(rng, (1,), dt)[0]
        y = ab.gemv(al, oa, A, x)                           # numpy operands: the host-buffer entry point
        ref = O.gemv(al, oa, A, x, wide=True)
        assert O.gemm_error_units(y, ref, A, x, x.shape[0], al) <= TOL_UNITS[dt], (O.NAMES[dt], oa)
        yd = host(ab.gemv(al, oa, dev(A, cuda), dev(x, cuda)))
        assert np.array_equal(y, yd) or O.gemm_error_units(y, yd, A, x, x.shape[0], al) <= TOL_UNITS[dt]
