"""CPU: pin the oracle (oracle/oracle.c) -- against the reference's haddock known answers, against
the committed OpenBLAS fixtures, and against a live OpenBLAS call with the reference's own CBLAS
arguments (LLVM/Native/Level3.hs:84-87, LLVM/Native/Level2.hs:67-70)."""
import numpy as np
import pytest

from oracle import oracle as O, openblas_ref as R
from _util import DT, OPN, as_np, gemm_inputs, gemv_inputs, haddock_cases, openblas_cases


@pytest.mark.parametrize("case", haddock_cases(), ids=lambda c: c["name"])
@pytest.mark.parametrize("wide", [False, True])
def test_haddock_known_answers(case, wide):
    dt = DT[case["dtype"]]
    exp = as_np(case["expect"], dt)
    if case["op"] == "dotu":
        got = O.dot(as_np(case["x"], dt), as_np(case["y"], dt), wide=wide)
    elif case["op"] == "outer":      # x >< y = (n,1) <> (1,n): LinearAlgebra.hs:73-78
        x, y = as_np(case["x"], dt), as_np(case["y"], dt)
        got = O.gemm(1, "N", x.reshape(-1, 1), "N", y.reshape(1, -1), wide=wide)
    elif case["op"] == "mXv":
        got = O.gemv(1, "N", as_np(case["m"], dt), as_np(case["x"], dt), wide=wide)
    elif case["op"] == "vXm":
        got = O.gemv(1, "T", as_np(case["m"], dt), as_np(case["x"], dt), wide=wide)
    else:
        got = O.gemm(1, "N", as_np(case["a"], dt), "N", as_np(case["b"], dt), wide=wide)
    assert np.array_equal(np.asarray(got), exp)     # small integers: exact


def test_openblas_fixture_gemm():
    z = openblas_cases()
    for i in range(int(z["n_gemm"][0])):
        dt, oa, ob = (int(v) for v in z[f"gemm{i}_meta"])
        A, B, al = z[f"gemm{i}_A"], z[f"gemm{i}_B"], z[f"gemm{i}_alpha"][0]
        k = A.shape[1] if oa == 0 else A.shape[0]
        for wide in (False, True):
            C = O.gemm(al, oa, A, ob, B, wide=wide)
            assert C.shape == z[f"gemm{i}_C"].shape
            assert O.gemm_error_units(z[f"gemm{i}_C"], C, A, B, k, al) < 1.0, (i, dt, oa, ob, wide)
            assert O.similar(z[f"gemm{i}_C"], C)    # the reference suite's own `~~~`


def test_openblas_fixture_gemv():
    z = openblas_cases()
    for i in range(int(z["n_gemv"][0])):
        dt, oa = (int(v) for v in z[f"gemv{i}_meta"])
        A, x, al = z[f"gemv{i}_A"], z[f"gemv{i}_x"], z[f"gemv{i}_alpha"][0]
        k = x.shape[0]
        for wide in (False, True):
            y = O.gemv(al, oa, A, x, wide=wide)
            assert O.gemm_error_units(z[f"gemv{i}_y"], y, A, x, k, al) < 1.0, (i, dt, oa, wide)
            assert O.similar(z[f"gemv{i}_y"], y)


@pytest.mark.parametrize("dt", [O.F32, O.F64, O.C32, O.C64], ids=lambda d: O.NAMES[d])
def test_live_openblas_matches_oracle(dt):
    rng = np.random.default_rng(100 + dt)
    for oa in OPN:
        for ob in OPN:
            m, n, k = (int(v) for v in rng.integers(1, 65, size=3))
            al, A, B = gemm_inputs(rng, dt, oa, ob, m, n, k)
            assert O.gemm_error_units(R.gemm(al, oa, A, ob, B), O.gemm(al, oa, A, ob, B), A, B, k, al) < 1.0
        m, n = (int(v) for v in rng.integers(1, 129, size=2))
        al, A, x = gemv_inputs(rng, dt, oa, m, n)
        assert O.gemm_error_units(R.gemv(al, oa, A, x), O.gemv(al, oa, A, x), A, x, x.shape[0], al) < 1.0
    x, y = O.uniform(rng, (1000,), dt), O.uniform(rng, (777,), dt)          # ragged: min extent
    d0, d1 = O.dot(x, y), R.dot(x[:777].copy(), y)
    assert abs(d0 - d1) <= 777 * O.EPS[dt] * np.linalg.norm(x[:777]) * np.linalg.norm(y)
    if dt in (O.C32, O.C64):
        d0, d1 = O.dot(x, y, conj=True), R.dot(x[:777].copy(), y, conj=True)
        assert abs(d0 - d1) <= 777 * O.EPS[dt] * np.linalg.norm(x[:777]) * np.linalg.norm(y)


def test_h_on_real_is_t():   # BLAS/Level3.hs:97-100
    rng = np.random.default_rng(3)
    al, A, B = gemm_inputs(rng, O.F64, "T", "T", 5, 6, 7)
    assert np.array_equal(O.gemm(al, "H", A, "H", B), O.gemm(al, "T", A, "T", B))


def test_edge_cases():
    z = np.zeros((3, 0))
    assert np.array_equal(O.gemm(1.0, "N", z, "N", np.zeros((0, 4))), np.zeros((3, 4)))   # k = 0: zeros
    assert O.gemm(1.0, "N", np.zeros((0, 5)), "N", np.zeros((5, 2))).shape == (0, 2)
    assert O.dot(np.zeros((0,)), np.zeros((0,))) == 0.0
    with pytest.raises(ValueError):
        O.gemm(1.0, "N", np.zeros((2, 3)), "N", np.zeros((4, 2)))
    with pytest.raises(ValueError):
        O.gemv(1.0, "N", np.zeros((2, 3)), np.zeros((2,)))
