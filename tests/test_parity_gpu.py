"""GPU: parity of the CUDA path (through the C ABI, via the python mirror) with the oracle.

Mirrors the reference's property tests (test/Test/BLAS/Level3.hs:31-72: gemm, dims 1..64, all nine
opA x opB, four dtypes, alpha and elements in [-1,1]; test/Test/BLAS/Level2.hs:31-68: gemv, dims
1..128) and adds what the reference omits: dot, zero extents, k = 1, n = 1,
unaligned / large shapes, the haddock golden vectors, the committed OpenBLAS fixtures.

Tolerances (floating point; stated here and in DESIGN.md):
  normwise  ||C - C_oracle||_F <= TOL * (k+2) * eps * |alpha| * ||A||_F * ||B||_F
  TOL = 1 (f32, f64), 2 (c32, c64: four real products per complex one), also on the 3xTF32 tensor path
  (measured there: 1e-4 .. 3e-4 units, the same as a sequential FP32 sum);
  plus the reference suite's own, much looser, elementwise `~~~` (abs 5e-5 or rel 5e-3).
"""
import numpy as np
import pytest

from oracle import oracle as O
from _util import (DT, OPN, TOL_UNITS, TOL_UNITS_TF32X3, as_np, gemm_inputs, gemv_inputs, haddock_cases,
                   openblas_cases)

pytestmark = pytest.mark.gpu
DTYPES = [O.F32, O.F64, O.C32, O.C64]
ids = lambda d: O.NAMES[d]  # noqa: E731


def dev(a, cuda):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(cuda)


def host(t):
    return t.cpu().numpy()


def tol_for(ab, dt, oa, ob, A, B):
    name = ab.kernel_name(dt, oa, ob, A.shape, B.shape)
    return TOL_UNITS_TF32X3 if "tf32x3" in name else TOL_UNITS[dt]   # tcgen05 and mma.sync 3xTF32 families


def check_gemm(ab, cuda, dt, oa, ob, al, A, B, path="auto"):
    ab.set_gemm_path(path)
    try:
        C = host(ab.gemm(al, oa, dev(A, cuda), ob, dev(B, cuda)))
        tol = tol_for(ab, dt, oa, ob, A, B)
    finally:
        ab.set_gemm_path("auto")
    k = A.shape[1] if oa == "N" else A.shape[0]
    ref = O.gemm(al, oa, A, ob, B, wide=True)
    assert C.shape == ref.shape and C.dtype == ref.dtype
    err = O.gemm_error_units(C, ref, A, B, k, al)
    assert err <= tol, (O.NAMES[dt], oa, ob, A.shape, B.shape, err)
    if k <= 128:   # the reference suite's own elementwise criterion, at the sizes it is used (k <= 64)
        assert O.similar(C, O.gemm(al, oa, A, ob, B, wide=False))
    return err


# --------------------------------------------------------------------------- golden vectors
@pytest.mark.parametrize("case", haddock_cases(), ids=lambda c: c["name"])
def test_haddock_known_answers(case, cuda):
    import accblas_b200 as ab
    for name in ("f32", "f64", "c32", "c64"):
        dt = DT[name]
        if case["dtype"] == "c64" and dt in (O.F32, O.F64):
            continue
        conv = (lambda v: as_np(v, DT[case["dtype"]]).astype(O.NP_DTYPE[dt]))
        exp = conv(case["expect"])
        if case["op"] == "dotu":
            got = host(ab.dot(dev(conv(case["x"]), cuda), dev(conv(case["y"]), cuda)))
        elif case["op"] == "outer":
            got = host(ab.outer(dev(conv(case["x"]), cuda), dev(conv(case["y"]), cuda)))
        elif case["op"] == "mXv":
            got = host(ab.mXv(dev(conv(case["m"]), cuda), dev(conv(case["x"]), cuda)))
        elif case["op"] == "vXm":
            got = host(ab.vXm(dev(conv(case["x"]), cuda), dev(conv(case["m"]), cuda)))
        else:
            got = host(ab.mXm(dev(conv(case["a"]), cuda), dev(conv(case["b"]), cuda)))
        assert np.array_equal(got, exp), (case["name"], name, got, exp)   # small integers: exact


def test_openblas_fixtures(cuda):
    import accblas_b200 as ab
    z = openblas_cases()
    for i in range(int(z["n_gemm"][0])):
        dt, oa, ob = (int(v) for v in z[f"gemm{i}_meta"])
        A, B, al = z[f"gemm{i}_A"], z[f"gemm{i}_B"], z[f"gemm{i}_alpha"][0]
        C = host(ab.gemm(al, oa, dev(A, cuda), ob, dev(B, cuda)))
        k = A.shape[1] if oa == 0 else A.shape[0]
        assert O.gemm_error_units(C, z[f"gemm{i}_C"], A, B, k, al) <= 2 * TOL_UNITS[dt], (i, dt, oa, ob)
    for i in range(int(z["n_gemv"][0])):
        dt, oa = (int(v) for v in z[f"gemv{i}_meta"])
        A, x, al = z[f"gemv{i}_A"], z[f"gemv{i}_x"], z[f"gemv{i}_alpha"][0]
        y = host(ab.gemv(al, oa, dev(A, cuda), dev(x, cuda)))
        assert O.gemm_error_units(y, z[f"gemv{i}_y"], A, x, x.shape[0], al) <= 2 * TOL_UNITS[dt], (i, dt, oa)


# --------------------------------------------------------------------------- gemm
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemm_property_small(dt, cuda):
    """The reference's generator: m, n, k in [1, 64], random op pair (Level3.hs:37-50, 68)."""
    import accblas_b200 as ab
    rng = np.random.default_rng(1000 + dt)
    for it in range(60):
        oa, ob = OPN[rng.integers(3)], OPN[rng.integers(3)]
        m, n, k = (int(v) for v in rng.integers(1, 65, size=3))
        al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
        check_gemm(ab, cuda, dt, oa, ob, al, A, B)


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
@pytest.mark.parametrize("oa", list(OPN))
@pytest.mark.parametrize("ob", list(OPN))
def test_gemm_all_ops_medium(dt, oa, ob, cuda):
    """Every op pair at tensor-path-eligible and at awkward extents."""
    import accblas_b200 as ab
    rng = np.random.default_rng(2000 + 16 * dt + 4 * OPN.index(oa) + OPN.index(ob))
    for (m, n, k) in [(256, 384, 320), (200, 136, 72), (129, 65, 33), (130, 258, 514)]:
        al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
        check_gemm(ab, cuda, dt, oa, ob, al, A, B)


@pytest.mark.parametrize("dt,knob,tiles", [(O.F64, 0, 4), (O.C64, 1, 3)], ids=["f64", "c64"])
def test_gemm_every_dmma_tile(dt, knob, tiles, cuda):
    """Each CTA tile shape of the DMMA family, forced through accblas_set_tuning, on every op pair at
    ragged extents (tile edges in m, n and k) and at one size with several tiles per dimension."""
    import accblas_b200 as ab
    rng = np.random.default_rng(2700 + dt)
    try:
        for tile in range(tiles):
            ab.set_tuning(knob, tile)
            for oa in OPN:
                for ob in OPN:
                    for (m, n, k) in [(200, 136, 72), (66, 34, 10), (300, 260, 130)]:
                        al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                        assert m < 128 or "dmma" in ab.kernel_name(dt, oa, ob, A.shape, B.shape)
                        check_gemm(ab, cuda, dt, oa, ob, al, A, B)
    finally:
        ab.set_tuning(knob, -1)


@pytest.mark.parametrize("dt", [O.F32, O.C32], ids=ids)
def test_gemm_mma_tf32_family(dt, cuda):
    """The one-launch mma.sync 3xTF32 family (f32 / c32, small and medium products), every CTA tile, every
    op pair, ragged extents, a k long enough for several accumulation chunks, windows of larger arrays."""
    import accblas_b200 as ab
    rng = np.random.default_rng(2800 + dt)
    try:
        ab.set_tuning(2, 1)                       # ACCBLAS_TUNE_SGEMM_FAMILY = mma.sync
        for tile in range(3):
            ab.set_tuning(3, tile)                # ACCBLAS_TUNE_SMMA_TILE
            for oa in OPN:
                for ob in OPN:
                    for (m, n, k) in [(200, 136, 72), (68, 36, 12), (300, 260, 1100)]:
                        al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                        assert "mma" in ab.kernel_name(dt, oa, ob, A.shape, B.shape)
                        C = host(ab.gemm(al, oa, dev(A, cuda), ob, dev(B, cuda)))
                        err = O.gemm_error_units(C, O.gemm(al, oa, A, ob, B, wide=True), A, B, k, al)
                        assert err <= TOL_UNITS_TF32X3, (O.NAMES[dt], tile, oa, ob, m, n, k, err)
                        if k <= 128:
                            assert O.similar(C, O.gemm(al, oa, A, ob, B, wide=False))
        ab.set_tuning(3, -1)
        # auto tile + windows (row pitch > columns)
        import torch
        for oa, ob in [("N", "N"), ("T", "H"), ("H", "T")]:
            al, A, B = gemm_inputs(rng, dt, oa, ob, 150, 90, 200)
            PA = dev(O.uniform(rng, (A.shape[0] + 2, A.shape[1] + 8), dt), cuda)
            PB = dev(O.uniform(rng, (B.shape[0] + 1, B.shape[1] + 4), dt), cuda)
            PA[1:1 + A.shape[0], 4:4 + A.shape[1]] = dev(A, cuda)
            PB[1:1 + B.shape[0], 4:4 + B.shape[1]] = dev(B, cuda)
            PC = torch.zeros((152, 96), dtype=PA.dtype, device=cuda)
            ab.gemm(al, oa, PA[1:1 + A.shape[0], 4:4 + A.shape[1]], ob, PB[1:1 + B.shape[0], 4:4 + B.shape[1]], out=PC[1:151, 4:94])
            C = np.ascontiguousarray(host(PC)[1:151, 4:94])
            assert O.gemm_error_units(C, O.gemm(al, oa, A, ob, B, wide=True), A, B, 200, al) <= TOL_UNITS_TF32X3
    finally:
        ab.set_tuning(2, -1)
        ab.set_tuning(3, -1)


@pytest.mark.parametrize("dt", [O.F32, O.C32], ids=ids)
def test_gemm_every_tc_tile(dt, cuda):
    """The tcgen05 3xTF32 family forced onto each tile (128 x 256 / 128 x 128 / 128 x 64 on one CTA, 256 x 256 on a CTA
    pair), every op pair, ragged extents in m, n and k (tile edges, odd pitches, an odd count of 128-row blocks for the
    pair), a k long enough for many promotion chunks.  Complex types cover the 4M embedding here; the 3M planes on the pair
    tile are covered by test_round2_gpu.py (BASELINE-size cgemm) and test_gemm_c3m_pair below."""
    import accblas_b200 as ab
    rng = np.random.default_rng(2900 + dt)
    try:
        ab.set_tuning(2, 2)                       # ACCBLAS_TUNE_SGEMM_FAMILY = tcgen05
        for tile in range(4):                     # 3 = the 256 x 256 CTA-pair tile (cta_group::2, round 2)
            ab.set_tuning(8, tile)                # ACCBLAS_TUNE_TC_TILE
            for oa in OPN:
                for ob in OPN:
                    for (m, n, k) in [(300, 260, 330), (129, 193, 1500)] + ([(257, 520, 200), (640, 300, 96)] if tile == 3 else []):
                        al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                        assert "tcgen05" in ab.kernel_name(dt, oa, ob, A.shape, B.shape)
                        C = host(ab.gemm(al, oa, dev(A, cuda), ob, dev(B, cuda)))
                        err = O.gemm_error_units(C, O.gemm(al, oa, A, ob, B, wide=True), A, B, k, al)
                        assert err <= TOL_UNITS_TF32X3, (O.NAMES[dt], tile, oa, ob, m, n, k, err)
    finally:
        ab.set_tuning(2, -1)
        ab.set_tuning(8, -1)


def test_gemm_c3m_pair(cuda):
    """cgemm 3M (three real products as stacked planes of ONE launch) on the CTA-pair tile: plane row offsets that are
    not multiples of 256, ragged extents; against the oracle."""
    import accblas_b200 as ab
    rng = np.random.default_rng(3131)
    try:
        ab.set_tuning(2, 2)
        ab.set_tuning(8, 3)
        ab.set_tuning(9, 1)                       # ACCBLAS_TUNE_CGEMM_3M
        for (m, n, k) in [(384, 260, 330), (1100, 1000, 520), (257, 300, 1000)]:
            for oa, ob in [("N", "N"), ("H", "N"), ("T", "H")]:
                al, A, B = gemm_inputs(rng, O.C32, oa, ob, m, n, k)
                assert "3m" in ab.kernel_name(O.C32, oa, ob, A.shape, B.shape)
                C = host(ab.gemm(al, oa, dev(A, cuda), ob, dev(B, cuda)))
                err = O.gemm_error_units(C, O.gemm(al, oa, A, ob, B, wide=True), A, B, k, al)
                assert err <= TOL_UNITS_TF32X3, (oa, ob, m, n, k, err)
    finally:
        ab.set_tuning(2, -1); ab.set_tuning(8, -1); ab.set_tuning(9, -1)


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemm_many_small_tiles_short_k(dt, cuda):
    """Hundreds of co-resident 32 x 32 CTAs whose smem ring wraps after a few K-blocks: the regime in
    which a consumer's stage release raced the producer's refill (fixed by the proxy fence in
    tma::consumer_release_fence).  The WHOLE result is compared, three times."""
    import accblas_b200 as ab
    rng = np.random.default_rng(2900 + dt)
    knob, tile = {O.F64: (0, 3), O.C64: (1, 2), O.F32: (3, 2), O.C32: (3, 2)}[dt]
    try:
        ab.set_tuning(knob, tile)
        if dt in (O.F32, O.C32):
            ab.set_tuning(2, 1)
        for (m, n, k) in [(768, 768, 96), (1024, 1024, 64)]:
            for oa, ob in [("N", "N"), ("T", "N"), ("N", "H")]:
                al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                ref = O.gemm(al, oa, A, ob, B, wide=True)
                dA, dB = dev(A, cuda), dev(B, cuda)
                for rep in range(3):
                    err = O.gemm_error_units(host(ab.gemm(al, oa, dA, ob, dB)), ref, A, B, k, al)
                    assert err <= TOL_UNITS_TF32X3, (O.NAMES[dt], oa, ob, m, n, k, rep, err)
    finally:
        ab.set_tuning(knob, -1)
        ab.set_tuning(2, -1)


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemm_generic_path_matches(dt, cuda):
    import accblas_b200 as ab
    rng = np.random.default_rng(2500 + dt)
    for oa in OPN:
        for ob in OPN:
            al, A, B = gemm_inputs(rng, dt, oa, ob, 96, 80, 112)
            check_gemm(ab, cuda, dt, oa, ob, al, A, B, path="generic")


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemm_degenerate(dt, cuda):
    """k = 1 (`><`, LinearAlgebra.hs:73-78), n = 1 (complex-H gemv route, PTX/Level2.hs:57-69),
    m = 1, zero extents (k = 0 must zero-fill an uninitialised C)."""
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(3000 + dt)
    for (m, n, k) in [(37, 41, 1), (64, 1, 50), (1, 70, 33), (1, 1, 1), (512, 512, 1), (1024, 1, 1024)]:
        for oa, ob in [("N", "N"), ("H", "N"), ("T", "H")]:
            al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
            check_gemm(ab, cuda, dt, oa, ob, al, A, B)
    tdt = torch.from_numpy(np.zeros(1, dtype=O.NP_DTYPE[dt])).dtype
    C = ab.gemm(1, "N", torch.zeros((5, 0), dtype=tdt, device=cuda), "N", torch.zeros((0, 7), dtype=tdt, device=cuda))
    assert C.shape == (5, 7) and torch.count_nonzero(C).item() == 0
    assert ab.gemm(1, "N", torch.zeros((0, 4), dtype=tdt, device=cuda), "N", torch.zeros((4, 3), dtype=tdt, device=cuda)).shape == (0, 3)
    with pytest.raises(ab.AccblasError):
        ab.gemm(1, "N", torch.zeros((2, 3), dtype=tdt, device=cuda), "N", torch.zeros((4, 2), dtype=tdt, device=cuda))


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemm_large_sampled(dt, cuda):
    """BASELINE-scale shapes: the oracle checks a sample of rows (it is O(mnk) scalar code); the
    full result is checked through linearity: (A1 + A2) B == A1 B + A2 B within tolerance."""
    import accblas_b200 as ab
    rng = np.random.default_rng(4000 + dt)
    n = 2048 if dt in (O.F32, O.F64) else 1024
    for oa, ob in [("N", "N"), ("N", "T"), ("T", "N"), ("H", "H")]:
        al, A, B = gemm_inputs(rng, dt, oa, ob, n, n, n)
        dA, dB = dev(A, cuda), dev(B, cuda)
        C = host(ab.gemm(al, oa, dA, ob, dB))
        tol = tol_for(ab, dt, oa, ob, A, B)
        rows = rng.choice(n, size=8, replace=False)
        Asub = A[rows] if oa == "N" else np.ascontiguousarray(A[:, rows])
        ref = O.gemm(al, oa, Asub, ob, B, wide=True)
        err = O.gemm_error_units(C[rows], ref, Asub, B, n, al)
        assert err <= tol, (O.NAMES[dt], oa, ob, err)
        A2 = O.uniform(rng, A.shape, dt)
        C2 = host(ab.gemm(al, oa, dev(A2, cuda), ob, dB))
        C12 = host(ab.gemm(al, oa, dev(A + A2, cuda), ob, dB))
        lin = O.gemm_error_units(C12, C + C2, A + A2, B, n, al)
        assert lin <= 3 * tol, (O.NAMES[dt], oa, ob, lin)


# --------------------------------------------------------------------------- gemv
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemv_property(dt, cuda):
    """The reference's generator: m, n in [1, 128], op in N/T/H (Level2.hs:37-46, 64) + larger."""
    import accblas_b200 as ab
    rng = np.random.default_rng(5000 + dt)
    shapes = [tuple(int(v) for v in rng.integers(1, 129, size=2)) for _ in range(40)]
    shapes += [(1, 1), (1, 300), (300, 1), (1000, 2000), (2000, 3000), (513, 1027), (4096, 64), (64, 4096),
               (3001, 2048), (2048, 3001)]
    for (m, n) in shapes:
        for oa in OPN:
            al, A, x = gemv_inputs(rng, dt, oa, m, n)
            y = host(ab.gemv(al, oa, dev(A, cuda), dev(x, cuda)))
            ref = O.gemv(al, oa, A, x, wide=True)
            assert y.shape == ref.shape
            err = O.gemm_error_units(y, ref, A, x, x.shape[0], al)
            assert err <= TOL_UNITS[dt], (O.NAMES[dt], oa, m, n, err)
            assert O.similar(y, O.gemv(al, oa, A, x, wide=False))


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemv_edges(dt, cuda):
    import torch
    import accblas_b200 as ab
    tdt = torch.from_numpy(np.zeros(1, dtype=O.NP_DTYPE[dt])).dtype
    y = ab.gemv(1, "N", torch.zeros((6, 0), dtype=tdt, device=cuda), torch.zeros((0,), dtype=tdt, device=cuda))
    assert y.shape == (6,) and torch.count_nonzero(y).item() == 0
    y = ab.gemv(1, "T", torch.zeros((0, 6), dtype=tdt, device=cuda), torch.zeros((0,), dtype=tdt, device=cuda))
    assert y.shape == (6,) and torch.count_nonzero(y).item() == 0
    assert ab.gemv(1, "N", torch.zeros((0, 6), dtype=tdt, device=cuda), torch.zeros((6,), dtype=tdt, device=cuda)).shape == (0,)
    with pytest.raises(ab.AccblasError):
        ab.gemv(1, "N", torch.zeros((3, 6), dtype=tdt, device=cuda), torch.zeros((3,), dtype=tdt, device=cuda))
    # unaligned base pointer (a view one element into a buffer)
    rng = np.random.default_rng(5500 + dt)
    al, A, x = gemv_inputs(rng, dt, "N", 67, 130)
    buf = torch.zeros(A.size + 1, dtype=tdt, device=cuda)
    buf[1:] = dev(A.reshape(-1), cuda)
    for oa in OPN:
        xx = x if oa == "N" else O.uniform(rng, (67,), dt)
        y = host(ab.gemv(al, oa, buf[1:].view(67, 130), dev(xx, cuda)))
        assert O.gemm_error_units(y, O.gemv(al, oa, A, xx), A, xx, xx.shape[0], al) <= TOL_UNITS[dt]


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemv_two_pass_path(dt, cuda):
    """Matrices above 64 MiB take the split-row two-pass T/H kernel (below: the one-launch slab kernel)."""
    import accblas_b200 as ab
    rng = np.random.default_rng(5700 + dt)
    m, n = 4200, 4100
    for oa in OPN:
        al, A, x = gemv_inputs(rng, dt, oa, m, n)
        y = host(ab.gemv(al, oa, dev(A, cuda), dev(x, cuda)))
        err = O.gemm_error_units(y, O.gemv(al, oa, A, x, wide=True), A, x, x.shape[0], al)
        assert err <= TOL_UNITS[dt], (O.NAMES[dt], oa, err)


def test_gemv_large_roundtrip(cuda):
    """BASELINE-size-independent property at a large shape: A (x1 + x2) == A x1 + A x2, and
    e_i^T A == row i, A e_j == column j (exact)."""
    import torch
    import accblas_b200 as ab
    m = n = 8192
    g = torch.Generator(device=cuda).manual_seed(7)
    A = torch.rand((m, n), dtype=torch.float64, device=cuda, generator=g) * 2 - 1
    e = torch.zeros(n, dtype=torch.float64, device=cuda)
    e[1234] = 1.0
    assert torch.equal(ab.mXv(A, e), A[:, 1234])
    assert torch.equal(ab.vXm(e, A), A[1234, :])
    x1 = torch.rand(n, dtype=torch.float64, device=cuda, generator=g) * 2 - 1
    x2 = torch.rand(n, dtype=torch.float64, device=cuda, generator=g) * 2 - 1
    for f in (lambda v: ab.mXv(A, v), lambda v: ab.vXm(v, A)):
        d = (f(x1 + x2) - (f(x1) + f(x2))).abs().max().item()
        assert d <= 3 * n * 2.0 ** -53 * 1.0 * (n ** 0.5), d


# --------------------------------------------------------------------------- dot
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_dot(dt, cuda):
    import accblas_b200 as ab
    rng = np.random.default_rng(6000 + dt)
    for n in [0, 1, 2, 3, 31, 32, 33, 1000, 2047, 2048, 2049, 65537, 1 << 20, (1 << 20) + 3, 5_000_001]:
        x, y = O.uniform(rng, (n,), dt), O.uniform(rng, (n,), dt)
        for conj in ([False, True] if dt in (O.C32, O.C64) else [False]):
            got = host((ab.dotc if conj else ab.dotu)(dev(x, cuda), dev(y, cuda)))
            ref = O.dot(x, y, conj=conj, wide=True)
            bound = (n + 2) * O.EPS[dt] * max(np.linalg.norm(x) * np.linalg.norm(y), 1e-300) * TOL_UNITS[dt]
            assert abs(complex(got) - complex(ref)) <= bound, (O.NAMES[dt], n, conj, got, ref)
    # ragged extents: min length (zipWith); unaligned views
    x, y = O.uniform(rng, (1000,), dt), O.uniform(rng, (777,), dt)
    got = host(ab.dotu(dev(x, cuda), dev(y, cuda)))
    assert abs(complex(got) - complex(O.dot(x, y))) <= 779 * O.EPS[dt] * np.linalg.norm(x) * np.linalg.norm(y) * 2
    dx, dy = dev(x, cuda), dev(x[::-1].copy(), cuda)
    got = host(ab.dotu(dx[1:], dy[3:]))
    ref = O.dot(x[1:], x[::-1][3:].copy())
    assert abs(complex(got) - complex(ref)) <= 1000 * O.EPS[dt] * np.linalg.norm(x) ** 2 * 2


def test_dot_is_reproducible_and_linear(cuda):
    import torch
    import accblas_b200 as ab
    g = torch.Generator(device=cuda).manual_seed(11)
    n = (1 << 26) + 5
    x = torch.rand(n, dtype=torch.float64, device=cuda, generator=g) * 2 - 1
    y = torch.rand(n, dtype=torch.float64, device=cuda, generator=g) * 2 - 1
    a, b = ab.dot(x, y), ab.dot(x, y)
    assert torch.equal(a, b)                                  # fixed summation order
    ones = torch.ones(n, dtype=torch.float64, device=cuda)
    assert ab.dot(ones, ones).item() == float(n)              # exact in fp64
    s = ab.dot(x, y).item() + ab.dot(x, ones).item()
    t = ab.dot(x, y + ones).item()
    assert abs(s - t) <= n * 2.0 ** -53 * (n ** 0.5) * 4


# --------------------------------------------------------------------------- host-buffer ABI
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_host_entry_points(dt, cuda):
    import accblas_b200 as ab
    rng = np.random.default_rng(7000 + dt)
    al, A, B = gemm_inputs(rng, dt, "N", "T", 300, 200, 100)
    C = ab.gemm(al, "N", A, "T", B)
    assert O.gemm_error_units(C, O.gemm(al, "N", A, "T", B), A, B, 100, al) <= TOL_UNITS_TF32X3
    al, A, B = gemm_inputs(rng, dt, "N", "N", 4096, 256, 128)        # row-panel pipeline path
    C = ab.gemm(al, "N", A, "N", B)
    ref = O.gemm(al, "N", A[::97], "N", B)
    assert O.gemm_error_units(C[::97], ref, A[::97], B, 128, al) <= TOL_UNITS_TF32X3
    al, A, x = gemv_inputs(rng, dt, "H", 150, 70)
    y = ab.gemv(al, "H", A, x)
    assert O.gemm_error_units(y, O.gemv(al, "H", A, x), A, x, 150, al) <= TOL_UNITS[dt]
    x, y = O.uniform(rng, (5000,), dt), O.uniform(rng, (5000,), dt)
    d = ab.dotu(x, y)
    assert abs(complex(d) - complex(O.dot(x, y))) <= 5002 * O.EPS[dt] * np.linalg.norm(x) * np.linalg.norm(y) * 2


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemm_strided_windows(dt, cuda):
    """accblas_gemm_ld: operands and result are windows of larger row-major arrays (row pitch >
    column count).  Every kernel family is reached: even pitches (TMA paths), odd pitches (SIMT),
    a product large enough for the tcgen05 path.  Elements of the parent of C outside the window
    must stay untouched."""
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(7500 + dt)
    for (m, n, k, pad, off) in [(256, 256, 256, 64, 32), (200, 136, 72, 8, 2), (130, 70, 50, 3, 1), (33, 17, 9, 5, 0)]:
        for oa in OPN:
            for ob in OPN:
                al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                PA = O.uniform(rng, (A.shape[0] + 3, A.shape[1] + pad), dt)
                PB = O.uniform(rng, (B.shape[0] + 2, B.shape[1] + pad), dt)
                PA[1:1 + A.shape[0], off:off + A.shape[1]] = A
                PB[2:2 + B.shape[0], off:off + B.shape[1]] = B
                dPA, dPB = dev(PA, cuda), dev(PB, cuda)
                sentinel = 7.0
                dPC = torch.full((m + 4, n + pad), sentinel, dtype=dPA.dtype, device=cuda)
                vA = dPA[1:1 + A.shape[0], off:off + A.shape[1]]
                vB = dPB[2:2 + B.shape[0], off:off + B.shape[1]]
                vC = dPC[3:3 + m, off:off + n]
                assert not vA.is_contiguous() and not vC.is_contiguous()
                ab.gemm(al, oa, vA, ob, vB, out=vC)
                PC = host(dPC)
                C = PC[3:3 + m, off:off + n]
                ref = O.gemm(al, oa, A, ob, B, wide=True)
                err = O.gemm_error_units(np.ascontiguousarray(C), ref, A, B, k, al)
                assert err <= TOL_UNITS_TF32X3, (O.NAMES[dt], oa, ob, m, n, k, err)
                PC[3:3 + m, off:off + n] = sentinel
                assert np.all(PC == sentinel), "gemm_ld wrote outside its window"


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_host_pipeline_wavefront(dt, cuda):
    """accblas_gemm_host on shapes that split into several row / column panels (csrc/host_pipeline.cu):
    P x Q block grids with P != Q, ragged last panels, every op pair; checked on sampled rows."""
    import accblas_b200 as ab
    rng = np.random.default_rng(7700 + dt)
    ab.set_tuning(4, 1)                           # ACCBLAS_TUNE_HOST_PIPELINE: wavefront whatever the size
    try:
        for (m, n, k) in [(1536, 1100, 96), (700, 2100, 64), (2304, 520, 40)]:
            for oa in OPN:
                for ob in OPN:
                    al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
                    C = ab.gemm(al, oa, A, ob, B)
                    rows = np.arange(0, m, 41)
                    Asub = A[rows] if oa == "N" else np.ascontiguousarray(A[:, rows])
                    ref = O.gemm(al, oa, Asub, ob, B, wide=True)
                    err = O.gemm_error_units(np.ascontiguousarray(C[rows]), ref, Asub, B, k, al)
                    assert err <= TOL_UNITS_TF32X3, (O.NAMES[dt], oa, ob, m, n, k, err)
    finally:
        ab.set_tuning(4, -1)
    # linearity over the whole result at a larger size (all blocks of the wavefront are covered)
    al, A, B = gemm_inputs(rng, dt, "N", "N", 2048, 2048, 256)
    A2 = O.uniform(rng, A.shape, dt)
    C1, C2, C12 = ab.gemm(al, "N", A, "N", B), ab.gemm(al, "N", A2, "N", B), ab.gemm(al, "N", A + A2, "N", B)
    assert O.gemm_error_units(C12, C1 + C2, A + A2, B, 256, al) <= 3 * TOL_UNITS_TF32X3


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_alpha_on_device(dt, cuda):
    """accblas_gemm_alpha_dev / accblas_gemv_alpha_dev: alpha read by the kernel from device memory gives
    bit-identical results to alpha passed by host pointer, in every kernel family."""
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(7900 + dt)
    shapes = [(40, 24, 1), (30, 20, 10), (200, 136, 72), (129, 65, 33), (512, 512, 512)]
    if dt in (O.F32, O.C32):
        shapes.append((1280, 1280, 1280))               # tcgen05 family
    for (m, n, k) in shapes:
        for oa, ob in [("N", "N"), ("H", "T")]:
            al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
            dA, dB = dev(A, cuda), dev(B, cuda)
            dal = dev(np.array([al], dtype=A.dtype), cuda)
            C0 = ab.gemm(al, oa, dA, ob, dB)
            C1 = ab.gemm(dal, oa, dA, ob, dB)
            assert torch.equal(C0, C1), (O.NAMES[dt], m, n, k, oa, ob, ab.kernel_name(dt, oa, ob, A.shape, B.shape))
    for (m, n) in [(67, 130), (1000, 2000), (4200, 4100)]:
        for oa in OPN:
            al, A, x = gemv_inputs(rng, dt, oa, m, n)
            dA, dx = dev(A, cuda), dev(x, cuda)
            dal = dev(np.array([al], dtype=A.dtype), cuda)
            assert torch.equal(ab.gemv(al, oa, dA, dx), ab.gemv(dal, oa, dA, dx)), (O.NAMES[dt], m, n, oa)


@pytest.mark.parametrize("force", [1, -1], ids=["forced", "auto"])
def test_zgemm_3m(force, cuda):
    """Complex-double gemm as three real DMMA products (csrc/gemm_3m.cu): every op pair, ragged and odd
    extents (plane pitch padding), alpha on the host and on the device, against the wide oracle under the
    stated complex tolerance; and the 4M kernel pinned by the knob gives the same product within it."""
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(9100 + force)
    shapes = [(200, 136, 72), (129, 65, 33), (301, 259, 131)] if force == 1 else [(2048, 2048, 2048)]
    try:
        ab.set_tuning(ab.TUNE_ZGEMM_3M, force)
        for (m, n, k) in shapes:
            for oa in OPN:
                for ob in (OPN if force == 1 else "NH"):
                    al, A, B = gemm_inputs(rng, O.C64, oa, ob, m, n, k)
                    assert ab.kernel_name(O.C64, oa, ob, A.shape, B.shape) == "zgemm_3m_dmma"
                    dA, dB = dev(A, cuda), dev(B, cuda)
                    C = host(ab.gemm(al, oa, dA, ob, dB))
                    rows = np.arange(m) if m <= 512 else rng.choice(m, size=8, replace=False)
                    Asub = A[rows] if oa == "N" else np.ascontiguousarray(A[:, rows])
                    ref = O.gemm(al, oa, Asub, ob, B, wide=True)
                    err = O.gemm_error_units(C[rows], ref, Asub, B, k, al)
                    assert err <= TOL_UNITS[O.C64], (oa, ob, m, n, k, err)
                    dal = dev(np.array([al], dtype=A.dtype), cuda)
                    assert torch.equal(ab.gemm(dal, oa, dA, ob, dB), torch.from_numpy(C).to(cuda))
                    ab.set_tuning(ab.TUNE_ZGEMM_3M, 0)
                    assert ab.kernel_name(O.C64, oa, ob, A.shape, B.shape) == "zgemm_dmma_tma"
                    C4 = host(ab.gemm(al, oa, dA, ob, dB))
                    ab.set_tuning(ab.TUNE_ZGEMM_3M, force)
                    assert O.gemm_error_units(C, C4, A, B, k, al) <= 2 * TOL_UNITS[O.C64]
    finally:
        ab.set_tuning(ab.TUNE_ZGEMM_3M, -1)


def test_cgemm_3m(cuda):
    """Complex-float gemm as three real tcgen05 3xTF32 products (pack_split3 / 3 x sgemm_tf32x3 / combine3):
    every op pair and CTA tile at ragged extents, alpha on host and device, against the wide oracle under the
    stated tolerance of the 3xTF32 path; 4M (real embedding) pinned by the knob agrees within it."""
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(9300)
    try:
        ab.set_tuning(2, 2)                       # tcgen05 family
        for tile, shapes in [(2, [(300, 260, 330), (129, 67, 1500)]), (1, [(257, 300, 200)]), (0, [(200, 520, 130)])]:
            ab.set_tuning(8, tile)
            for (m, n, k) in shapes:
                for oa in OPN:
                    for ob in OPN:
                        al, A, B = gemm_inputs(rng, O.C32, oa, ob, m, n, k)
                        dA, dB = dev(A, cuda), dev(B, cuda)
                        ab.set_tuning(ab.TUNE_CGEMM_3M, 1)
                        assert ab.kernel_name(O.C32, oa, ob, A.shape, B.shape) == "cgemm_3m_tf32x3_tcgen05"
                        C = host(ab.gemm(al, oa, dA, ob, dB))
                        dal = dev(np.array([al], dtype=A.dtype), cuda)
                        assert torch.equal(ab.gemm(dal, oa, dA, ob, dB), torch.from_numpy(C).to(cuda))
                        ab.set_tuning(ab.TUNE_CGEMM_3M, 0)
                        assert ab.kernel_name(O.C32, oa, ob, A.shape, B.shape) == "cgemm_tf32x3_tcgen05"
                        C4 = host(ab.gemm(al, oa, dA, ob, dB))
                        err = O.gemm_error_units(C, O.gemm(al, oa, A, ob, B, wide=True), A, B, k, al)
                        assert err <= TOL_UNITS_TF32X3, (tile, oa, ob, m, n, k, err)
                        assert O.gemm_error_units(C, C4, A, B, k, al) <= 2 * TOL_UNITS_TF32X3
    finally:
        for knob in (2, 8, ab.TUNE_CGEMM_3M):
            ab.set_tuning(knob, -1)


def test_dgemm_streamk(cuda):
    """The persistent stream-K variant of the 128 x 128 DMMA kernel (forced): tiles cut across several CTAs
    (few K-blocks per CTA), whole tiles plus head / tail fixups, ragged edges, every op pair; against the wide
    oracle on sampled rows and against the data-parallel kernel on the whole matrix; run-to-run bit equality."""
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(9500)
    try:
        for (m, n, k) in [(640, 512, 512), (1000, 900, 777), (300, 260, 4000), (2048, 2048, 640)]:
            for oa, ob in [("N", "N"), ("T", "N"), ("N", "T"), ("T", "T")]:
                al, A, B = gemm_inputs(rng, O.F64, oa, ob, m, n, k)
                dA, dB = dev(A, cuda), dev(B, cuda)
                ab.set_tuning(ab.TUNE_STREAMK, 1)
                C1 = ab.gemm(al, oa, dA, ob, dB)
                C1b = ab.gemm(al, oa, dA, ob, dB)
                ab.set_tuning(ab.TUNE_STREAMK, 0)
                C0 = ab.gemm(al, oa, dA, ob, dB)
                assert torch.equal(C1, C1b), (m, n, k, oa, ob)
                C = host(C1)
                assert O.gemm_error_units(C, host(C0), A, B, k, al) <= 2 * TOL_UNITS[O.F64], (m, n, k, oa, ob)
                rows = rng.choice(m, size=6, replace=False)
                Asub = A[rows] if oa == "N" else np.ascontiguousarray(A[:, rows])
                err = O.gemm_error_units(C[rows], O.gemm(al, oa, Asub, ob, B, wide=True), Asub, B, k, al)
                assert err <= TOL_UNITS[O.F64], (m, n, k, oa, ob, err)
    finally:
        ab.set_tuning(ab.TUNE_STREAMK, -1)


def test_cpp_host_mirror(cuda):
    """The C++ mirror of the reference's interface (include/accblas_b200.hpp) through its own test program:
    haddock known answers, random gemm / gemv / dot over 4 dtypes x N/T/H, Level-1 extras, error behaviour."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "cpp", "hpp_check")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-s", "-C", os.path.dirname(exe)])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "hpp_check ok" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gemm_strided_batched(dt, cuda):
    """accblas_gemm_strided_batched: tiny products (one z-batched SIMT launch), products of the warp-level tensor families
    (ONE launch: the batch is a grid dimension and the third dimension of the TMA maps -- round 2), strides that are not
    16-byte multiples (loop of launches), a broadcast operand (stride 0) on either side; every product against the oracle."""
    import torch
    import accblas_b200 as ab
    rng = np.random.default_rng(9700 + dt)
    es = {O.F32: 4, O.F64: 8, O.C32: 8, O.C64: 16}[dt]
    cases = [(7, 5, 9, 13, "N", "N"), (3, 33, 20, 17, "H", "T"), (4, 136, 72, 200, "T", "N"), (16, 100, 100, 100, "N", "N"),
             (5, 96, 160, 64, "N", "H"), (3, 250, 130, 75, "H", "N"), (6, 64, 66, 40, "T", "T"), (2, 520, 384, 96, "N", "T")]
    one_launch = 0
    for (batch, m, n, k, oa, ob) in cases:
        As = [gemm_inputs(rng, dt, oa, ob, m, n, k) for _ in range(batch)]
        al = As[0][0]
        A = np.stack([a for (_, a, _) in As])
        B = np.stack([b for (_, _, b) in As])
        dA, dB = dev(A, cuda), dev(B, cuda)
        name = ab.kernel_name(dt, oa, ob, A[0].shape, B[0].shape)
        before = ab.launch_count()
        C = host(ab.gemm_batched(al, oa, dA, ob, dB))
        launches = ab.launch_count() - before
        strides16 = all((es * x) % 16 == 0 for x in (A[0].size, B[0].size, m * n, A.shape[-1], B.shape[-1]))   # planes and row pitches
        if m * n * k < 48 ** 3 and "generic" in name:
            assert launches == 1, (launches, m, n, k)
        elif strides16 and ("dmma_tma" in name or "_mma" in name) and "3m" not in name:
            assert launches == 1, (name, launches, m, n, k)
            one_launch += 1
        else:
            assert launches >= batch, (name, launches, m, n, k)
        for b in range(batch):
            err = O.gemm_error_units(C[b], O.gemm(al, oa, A[b], ob, B[b], wide=True), A[b], B[b], k, al)
            assert err <= TOL_UNITS_TF32X3, (O.NAMES[dt], name, b, m, n, k, err)
        # B, then A, broadcast over the batch (stride 0)
        Cb = host(ab.gemm_batched(al, oa, dA, ob, dev(B[0], cuda)))
        for b in range(batch):
            assert O.gemm_error_units(Cb[b], O.gemm(al, oa, A[b], ob, B[0], wide=True), A[b], B[0], k, al) <= TOL_UNITS_TF32X3
        Ca = host(ab.gemm_batched(al, oa, dev(A[0], cuda), ob, dB))
        for b in range(batch):
            assert O.gemm_error_units(Ca[b], O.gemm(al, oa, A[0], ob, B[b], wide=True), A[0], B[b], k, al) <= TOL_UNITS_TF32X3
    assert one_launch >= 3, one_launch
    with pytest.raises(ab.AccblasError):
        ab.gemm_batched(1, "N", torch.zeros((2, 3, 4), dtype=dev(A, cuda).dtype, device=cuda), "N",
                        torch.zeros((2, 5, 2), dtype=dev(A, cuda).dtype, device=cuda))


def test_launches_are_counted(cuda):
    import torch
    import accblas_b200 as ab
    a = torch.ones((64, 64), dtype=torch.float64, device=cuda)
    before = ab.launch_count()
    ab.mXm(a, a)
    assert ab.launch_count() > before
