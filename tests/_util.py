"""Shared helpers for the test-suite (test infrastructure: may import oracle/)."""
import json
import os

import numpy as np

from oracle import oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DT = {"f32": O.F32, "f64": O.F64, "c32": O.C32, "c64": O.C64}
OPN = "NTH"


def haddock_cases():
    with open(os.path.join(GOLDEN, "haddock.json")) as f:
        return json.load(f)


def as_np(v, dt):
    a = np.array(v, dtype=np.float64)
    if dt in (O.C32, O.C64):
        a = a[..., 0] + 1j * a[..., 1]
    return a.astype(O.NP_DTYPE[dt])


def openblas_cases():
    return np.load(os.path.join(GOLDEN, "openblas_cases.npz"))


def gemm_inputs(rng, dt, oa, ob, m, n, k):
    """Stored operands have the PRE-op shape (test/Test/BLAS/Level3.hs:45-50)."""
    A = O.uniform(rng, (m, k) if oa == "N" else (k, m), dt)
    B = O.uniform(rng, (k, n) if ob == "N" else (n, k), dt)
    alpha = O.uniform(rng, (1,), dt)[0]
    return alpha, A, B


def gemv_inputs(rng, dt, oa, m, n):
    A = O.uniform(rng, (m, n), dt)
    x = O.uniform(rng, (n if oa == "N" else m,), dt)
    alpha = O.uniform(rng, (1,), dt)[0]
    return alpha, A, x


# Stated tolerances (DESIGN.md "Parity"): normwise error in units of k * eps * |alpha| * ||A||_F * ||B||_F
TOL_UNITS = {O.F32: 1.0, O.F64: 1.0, O.C32: 2.0, O.C64: 2.0}
# the tcgen05 3xTF32 path carries a 2^-22 operand-split error per product; with the chunked FP32
# promotion of the accumulator it measures 1e-4 .. 3e-4 units, so it gets the same bound as complex
TOL_UNITS_TF32X3 = 2.0
