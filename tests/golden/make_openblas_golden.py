"""Writes tests/golden/openblas_cases.npz: seeded inputs and the outputs of the library the
reference's llvm-native hooks call (cblas_?gemm / ?gemv RowMajor, LLVM/Native/Level3.hs:84-87,
LLVM/Native/Level2.hs:67-70), produced HERE by the scipy-bundled OpenBLAS through
oracle/openblas_ref.py.  The fixture pins the oracle (tests/test_oracle.py) and the CUDA path
(tests/test_parity_gpu.py) to that library's results without needing it at test time.
Run from the repo root:  python tests/golden/make_openblas_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O, openblas_ref as R  # noqa: E402

rng = np.random.default_rng(0x5EED)
out = {}
idx = 0
for dt in (O.F32, O.F64, O.C32, O.C64):
    for oa in "NTH":
        for ob in "NTH":
            m, n, k = (int(v) for v in rng.integers(1, 65, size=3))     # test/Test/BLAS/Level3.hs:68
            A = O.uniform(rng, (m, k) if oa == "N" else (k, m), dt)
            B = O.uniform(rng, (k, n) if ob == "N" else (n, k), dt)
            al = O.uniform(rng, (1,), dt)[0]
            out[f"gemm{idx}_meta"] = np.array([dt, "NTH".index(oa), "NTH".index(ob)])
            out[f"gemm{idx}_alpha"] = np.array([al])
            out[f"gemm{idx}_A"], out[f"gemm{idx}_B"] = A, B
            out[f"gemm{idx}_C"] = R.gemm(al, oa, A, ob, B)
            idx += 1
out["n_gemm"] = np.array([idx])
idx = 0
for dt in (O.F32, O.F64, O.C32, O.C64):
    for oa in "NTH":
        m, n = (int(v) for v in rng.integers(1, 129, size=2))           # test/Test/BLAS/Level2.hs:64
        A = O.uniform(rng, (m, n), dt)
        x = O.uniform(rng, (n if oa == "N" else m,), dt)
        al = O.uniform(rng, (1,), dt)[0]
        out[f"gemv{idx}_meta"] = np.array([dt, "NTH".index(oa)])
        out[f"gemv{idx}_alpha"] = np.array([al])
        out[f"gemv{idx}_A"], out[f"gemv{idx}_x"] = A, x
        out[f"gemv{idx}_y"] = R.gemv(al, oa, A, x)
        idx += 1
out["n_gemv"] = np.array([idx])
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "openblas_cases.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path), "bytes")
