"""sdot / asum / amax / amin (BLAS/Level1.hs:44-62, 100-143) and the outer product `><`
(LinearAlgebra.hs:73-78).

The reference has no test and no haddock example for the four reductions (test/Test/BLAS.hs:29-32 runs
Level 2 and 3 only), so their parity is pinned (CPU half of this file) on
  * the definitions themselves, restated a second time with numpy (independent of oracle.c), and
  * the CBLAS routines the reference's haddocks name as the model (cblas_sdot family / ?asum / i?amax),
    reached in the scipy-bundled OpenBLAS -- equal up to the documented tie rule: the reference returns the
    LAST index among equal extrema (`cmp ix iy = snd ix > snd iy ? ix : iy`), CBLAS the first.
The GPU half compares the CUDA kernels with the oracle: indices bit-exact, sums within
(n + 2) * eps * sum|x_i y_i| (eps of the OUTPUT type; sdot's double accumulation leaves only the final rounding).
"""
import numpy as np
import pytest

from oracle import oracle as O

DTYPES = [O.F32, O.F64, O.C32, O.C64]
ids = lambda d: O.NAMES[d]  # noqa: E731
SIZES = [1, 2, 3, 31, 32, 33, 100, 511, 512, 513, 2047, 4096, 4097, 65537, (1 << 20) + 3]


def mags(x):
    """`abs` / `mag c = abs (real c) + abs (imag c)` in the element's own precision."""
    if np.iscomplexobj(x):
        return np.abs(x.real) + np.abs(x.imag)
    return np.abs(x)


def last_arg(m, want_max):
    best = m.max() if want_max else m.min()
    return int(np.flatnonzero(m == best)[-1])


# ------------------------------------------------------------------------------- CPU: oracle pins
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_oracle_matches_numpy_restatement(dt):
    rng = np.random.default_rng(8000 + dt)
    for n in SIZES[:12]:
        x, y = O.uniform(rng, (n,), dt), O.uniform(rng, (n,), dt)
        z = O.uniform(rng, (1,), dt)[0]
        wide = np.complex128 if dt in (O.C32, O.C64) else np.float64
        ref = (wide(z) + np.sum(x.astype(wide) * y.astype(wide))).astype(O.NP_DTYPE[dt])
        for w in (False, True):
            got = O.sdot(z, x, y, wide=w)
            assert abs(complex(got) - complex(ref)) <= 4 * O.EPS[dt] * (abs(z) + np.sum(np.abs(x) * np.abs(y))) + 1e-300
        base = np.float32 if dt in (O.F32, O.C32) else np.float64
        s = O.asum(x, wide=True)
        assert s.dtype == base
        assert abs(float(s) - float(np.sum(mags(x).astype(np.float64)))) <= 2 * O.EPS[dt] * float(np.sum(mags(x)))
        assert abs(float(O.asum(x, wide=False)) - float(s)) <= (n + 2) * O.EPS[dt] * float(s)
        assert O.amax(x) == last_arg(mags(x), True)
        assert O.amin(x) == last_arg(mags(x), False)


@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_oracle_tie_rule_is_last_index(dt):
    """`cmp ix iy = snd ix > snd iy ? ix : iy`: equal magnitudes keep the RIGHT operand."""
    x = np.array([1, -3, 2, 3, -1, 0.5, -3, 1], dtype=O.NP_DTYPE[dt])
    assert O.amax(x) == 6 and O.amin(x) == 5
    x = np.array([2, -2, 2, 2], dtype=O.NP_DTYPE[dt])
    assert O.amax(x) == 3 and O.amin(x) == 3
    if dt in (O.C32, O.C64):
        z = np.array([1 + 1j, -2 + 0j, 0 - 2j, 0.5 + 0.5j, -1 + 0j], dtype=O.NP_DTYPE[dt])
        assert O.amax(z) == 2 and O.amin(z) == 4          # |re| + |im|: 2, 2, 2, 1, 1
    with pytest.raises(ValueError):
        O.amax(np.zeros((0,), dtype=O.NP_DTYPE[dt]))      # fold1 over an empty array


def test_oracle_matches_openblas_level1():
    import scipy.linalg.blas as sb
    rng = np.random.default_rng(8100)
    for n in (1, 7, 100, 4097):
        for dt, asum_f, amax_f in ((O.F32, sb.sasum, sb.isamax), (O.F64, sb.dasum, sb.idamax),
                                   (O.C32, sb.scasum, sb.icamax), (O.C64, sb.dzasum, sb.izamax)):
            x = O.uniform(rng, (n,), dt)
            assert abs(float(O.asum(x)) - float(asum_f(x))) <= (n + 2) * O.EPS[dt] * float(O.asum(x))
            m = mags(x)
            if np.count_nonzero(m == m.max()) == 1:       # no tie: first == last
                assert O.amax(x) == int(amax_f(x))
    # cblas_sdsdot / dsdot = "sdot with double accumulation" (the haddock's model for sdot)
    from oracle import openblas_ref as R
    L = R.lib()
    import ctypes
    L.scipy_cblas_sdsdot.restype = ctypes.c_float
    L.scipy_cblas_sdsdot.argtypes = [ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    x, y = O.uniform(rng, (5001,), O.F32), O.uniform(rng, (5001,), O.F32)
    ref = L.scipy_cblas_sdsdot(5001, ctypes.c_float(0.25), x.ctypes.data, 1, y.ctypes.data, 1)
    got = O.sdot(np.float32(0.25), x, y)
    assert abs(float(got) - float(ref)) <= 2 * 2.0 ** -24 * abs(float(ref)) + 1e-7


# ------------------------------------------------------------------------------- GPU: parity
def dev(a, cuda):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(cuda)


@pytest.mark.gpu
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gpu_level1_extras(dt, cuda):
    import accblas_b200 as ab
    rng = np.random.default_rng(8200 + dt)
    base = np.float32 if dt in (O.F32, O.C32) else np.float64
    for n in SIZES + [5_000_001]:
        x, y = O.uniform(rng, (n,), dt), O.uniform(rng, (n,), dt)
        z = O.uniform(rng, (1,), dt)[0]
        dx, dy = dev(x, cuda), dev(y, cuda)
        got = ab.sdot(z, dx, dy).cpu().numpy()
        ref = O.sdot(z, x, y, wide=True)
        scale = abs(z) + float(np.sum(np.abs(x).astype(np.float64) * np.abs(y).astype(np.float64)))
        # double accumulation: error = (n+2) * 2^-53 * scale from the sum + one rounding to the element type
        assert abs(complex(got) - complex(ref)) <= (n + 2) * 2.0 ** -53 * scale + 2 * O.EPS[dt] * abs(complex(ref)) + 1e-300, (n, got, ref)
        s = ab.asum(dx).cpu().numpy()
        assert s.dtype == base
        sref = float(O.asum(x, wide=True))
        assert abs(float(s) - sref) <= 2 * O.EPS[dt] * sref + (n + 2) * 2.0 ** -53 * sref, (n, s, sref)
        assert int(ab.amax(dx).item()) == O.amax(x), n       # index work: bit-exact
        assert int(ab.amin(dx).item()) == O.amin(x), n
    # ties across CTA / warp / vector boundaries: the LAST index must win
    n = 3_000_000
    x = O.uniform(rng, (n,), dt) * 0.5
    big = np.array([3.0], dtype=O.NP_DTYPE[dt])[0]
    for pos in ([0, 1], [5, n - 1], [4095, 4096, 2_000_000], [n - 3, n - 2]):
        xx = x.copy()
        xx[pos] = big
        xx[[p + 7 for p in pos if p + 7 < n]] = 0            # and ties for the minimum
        d = dev(xx, cuda)
        assert int(ab.amax(d).item()) == O.amax(xx) == max(pos)
        assert int(ab.amin(d).item()) == O.amin(xx)
    # unaligned views, ragged sdot extents (zipWith), host-buffer entry points, error statuses
    dxx = dev(x, cuda)
    assert int(ab.amax(dxx[1:100_001]).item()) == O.amax(x[1:100_001])
    assert abs(float(ab.asum(dxx[3:777]).item()) - float(O.asum(x[3:777]))) <= 800 * O.EPS[dt] * float(O.asum(x[3:777]))
    g = ab.sdot(z, dxx[1:1001], dxx[2:900]).cpu().numpy()
    r = O.sdot(z, x[1:1001], x[2:900])
    assert abs(complex(g) - complex(r)) <= 1000 * O.EPS[dt] * (abs(z) + 900)
    xs = x[:70_001]
    assert ab.amax(xs) == O.amax(xs) and ab.amin(xs) == O.amin(xs)
    assert abs(float(ab.asum(xs)) - float(O.asum(xs))) <= 4 * O.EPS[dt] * float(O.asum(xs))
    assert abs(complex(ab.sdot(z, xs, xs[::-1].copy())) - complex(O.sdot(z, xs, xs[::-1].copy()))) <= 4 * O.EPS[dt] * (abs(z) + 70_001)
    with pytest.raises(ab.AccblasError):
        ab.amax(dxx[:0])
    assert float(ab.asum(dxx[:0]).item()) == 0.0
    assert complex(ab.sdot(z, dxx[:0], dxx[:0]).cpu().numpy()) == complex(z)


@pytest.mark.gpu
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gpu_outer_product_rank1(dt, cuda):
    """`x >< y` and every other k = 1 product go through the rank-1 store kernel: exact for real
    types against the literal fold ((alpha * a) * b, one product, no sum)."""
    import accblas_b200 as ab
    rng = np.random.default_rng(8300 + dt)
    for (m, n) in [(3, 3), (1, 1), (37, 41), (1000, 2000), (2049, 513), (5, 70001)]:
        x, y = O.uniform(rng, (m,), dt), O.uniform(rng, (n,), dt)
        assert ab.kernel_name(dt, "N", "N", (m, 1), (1, n)) == "gemm_rank1_outer"
        got = ab.outer(dev(x, cuda), dev(y, cuda)).cpu().numpy()
        ref = O.gemm(1, "N", x.reshape(-1, 1), "N", y.reshape(1, -1), wide=False)
        if dt in (O.F32, O.F64):
            assert np.array_equal(got, ref)
        else:
            assert O.gemm_error_units(got, ref, x, y, 1) <= 1.0
        for oa, ob in [("T", "N"), ("H", "H"), ("N", "T"), ("H", "N")]:
            al = O.uniform(rng, (1,), dt)[0]
            A = x.reshape(-1, 1) if oa == "N" else x.reshape(1, -1)
            B = y.reshape(1, -1) if ob == "N" else y.reshape(-1, 1)
            got = ab.gemm(al, oa, dev(A, cuda), ob, dev(B, cuda)).cpu().numpy()
            ref = O.gemm(al, oa, A, ob, B, wide=True)
            assert O.gemm_error_units(got, ref, A, B, 1, al) <= 1.0, (oa, ob, m, n)


# ------------------------------------------------------------------------------- compensated dot
# The reference's only known-answer vector for its compensated sums (Numeric/Sum.hs:72-80):
#   sum kbn [1.0, 1.0e100, 1.0, -1.0e100] == 2.0      (Prelude.sum gives 0.0)
KBN_HADDOCK = ([1.0, 1.0e100, 1.0, -1.0e100], 2.0)


def illcond(rng, n, dt, spread):
    """x . y with heavy cancellation: the second half undoes the first up to a small perturbation, exponents spread
    over 2^+-spread.  Returns operands whose exact dot is ~1e-10 .. 1e-6 of sum |x_i y_i|."""
    h = n // 2
    base = np.float32 if dt in (O.F32, O.C32) else np.float64

    def part(shape):
        return (rng.uniform(-1, 1, shape) * 2.0 ** rng.integers(-spread, spread, shape)).astype(base)
    if dt in (O.C32, O.C64):
        a = part(h) + 1j * part(h)
        b = part(h) + 1j * part(h)
    else:
        a, b = part(h), part(h)
    a = a.astype(O.NP_DTYPE[dt])
    b = b.astype(O.NP_DTYPE[dt])
    x = np.concatenate([a, a])
    y = np.concatenate([b, -b])
    y[h:] += (np.abs(b) * 1e-6).astype(y.dtype) * (rng.uniform(-1, 1, h)).astype(base)
    p = rng.permutation(2 * h)
    return np.ascontiguousarray(x[p]), np.ascontiguousarray(y[p])


def test_oracle_kbn_restatement_and_exact_dot():
    xs, want = KBN_HADDOCK
    assert O.kbn_sum(np.array(xs, dtype=np.float64)) == want          # the reference's own example
    assert float(np.sum(np.array(xs))) != want                        # ... which a plain sum misses
    rng = np.random.default_rng(8300)
    x, y = illcond(rng, 2000, O.F64, 30)
    import math
    assert abs(O.dot_exact(x, y)) <= 1e-3 * float(np.sum(np.abs(x) * np.abs(y)))      # the generator does cancel
    # products of floats are exact in double, so fsum of them is the exact dot rounded once: a second route to dot_exact
    xf, yf = illcond(rng, 2000, O.F32, 10)
    assert O.dot_exact(xf, yf) == math.fsum(float(a) * float(b) for a, b in zip(xf, yf))


@pytest.mark.gpu
@pytest.mark.parametrize("dt", DTYPES, ids=ids)
def test_gpu_dot_compensated(dt, cuda):
    import accblas_b200 as ab
    cplx = dt in (O.C32, O.C64)
    # the reference's haddock vector, as a dot with ones (element types that can hold 1e100)
    if dt in (O.F64, O.C64):
        xs, want = KBN_HADDOCK
        x = np.array(xs, dtype=O.NP_DTYPE[dt])
        one = np.ones(4, dtype=O.NP_DTYPE[dt])
        assert complex(ab.dot_compensated(dev(x, cuda), dev(one, cuda)).item()) == complex(want)
        if dt == O.F64:      # the plain fold loses the ones (for c64 its tree happens to pair the big terms first)
            assert complex(ab.dotu(dev(x, cuda), dev(one, cuda)).item()) != complex(want)
    rng = np.random.default_rng(8400 + dt)
    assert complex(ab.dot_compensated(dev(np.zeros(0, O.NP_DTYPE[dt]), cuda), dev(np.zeros(0, O.NP_DTYPE[dt]), cuda)).item()) == 0
    for n in (2, 3, 33, 1000, 4097, 65538, 200_001):
        for conj in ((False, True) if cplx else (False,)):
            x, y = illcond(rng, n, dt, 8 if dt in (O.F32, O.C32) else 30)
            x, y = x[: n], y[: max(1, n - 1)]                          # unequal extents: zipWith intersects
            exact = complex(O.dot_exact(x, y, conj))
            got = complex(ab.dot_compensated(dev(x, cuda), dev(y, cuda), conj).item())
            plain = complex((ab.dotc if conj else ab.dotu)(dev(x, cuda), dev(y, cuda)).item())
            mag = float(np.sum(np.abs(x[: len(y)]).astype(np.float64) * np.abs(y).astype(np.float64)))
            # twice-double-precision evaluation + one rounding to the element type
            bound = 2 * O.EPS[dt] * abs(exact) + (2 if cplx else 1) * (len(y) * 2.0 ** -52) ** 2 * mag + 1e-300
            assert abs(got - exact) <= bound, (O.NAMES[dt], n, conj, got, exact, bound)
            # the plain fold is allowed its usual n * eps * sum|x_i y_i| -- and on these inputs it needs it
            assert abs(plain - exact) <= (len(y) + 2) * O.EPS[dt] * mag
    # unaligned operands take the scalar path
    x, y = illcond(rng, 10_001, dt, 8 if dt in (O.F32, O.C32) else 30)
    dx, dy = dev(x, cuda), dev(y, cuda)
    got = complex(ab.dot_compensated(dx[1:], dy[1:]).item())
    exact = complex(O.dot_exact(x[1:], y[1:]))
    mag = float(np.sum(np.abs(x[1:]).astype(np.float64) * np.abs(y[1:]).astype(np.float64)))
    assert abs(got - exact) <= 2 * O.EPS[dt] * abs(exact) + 2 * (10_000 * 2.0 ** -52) ** 2 * mag + 1e-300
    # run-to-run bit equality (fixed tree)
    a = ab.dot_compensated(dx, dy).cpu().numpy().tobytes()
    assert all(ab.dot_compensated(dx, dy).cpu().numpy().tobytes() == a for _ in range(5))
