"""tools/exp/c3m_sweep.py -- c32 gemm: 3M (three real tcgen05 products) against the 4M real embedding."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import perf  # noqa: E402
import accblas_b200 as ab  # noqa: E402

for n in (768, 1000, 1024, 1536, 2048, 3072, 4096, 8192):
    A, B = perf.rnd((n, n), "c32", 1), perf.rnd((n, n), "c32", 2)
    row = {"sweep": "c32gemm_HN", "n": n}
    for name, v in (("m4", 0), ("m3", 1), ("auto", -1)):
        ab.set_tuning(ab.TUNE_CGEMM_3M, v)
        med, best = perf.timeit(lambda: ab.gemm(1, "H", A, "N", B), 20 if n <= 4096 else 5)
        row[name + "_us"] = round(med * 1e3, 1)
        row[name + "_kernel"] = ab.kernel_name(2, "H", "N", A.shape, B.shape)
    ab.set_tuning(ab.TUNE_CGEMM_3M, -1)
    row["tflops_3m"] = round(8.0 * n ** 3 / row["m3_us"] * 1e-6, 1)
    print(json.dumps(row), flush=True)
