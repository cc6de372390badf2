"""tools/exp/streamk_sweep.py -- f64 gemm: persistent stream-K kernel against the data-parallel tile family."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import perf  # noqa: E402
import accblas_b200 as ab  # noqa: E402

for n in (768, 1000, 1024, 1280, 1536, 2048, 2560, 3072, 4096, 8192):
    A, B = perf.rnd((n, n), "f64", 1), perf.rnd((n, n), "f64", 2)
    row = {"sweep": "f64gemm_NN", "n": n}
    for name, v in (("dp", 0), ("streamk", 1), ("auto", -1)):
        ab.set_tuning(ab.TUNE_STREAMK, v)
        med, best = perf.timeit(lambda: ab.gemm(1, "N", A, "N", B), 30 if n <= 2048 else 5)
        row[name + "_us"] = round(med * 1e3, 1)
    ab.set_tuning(ab.TUNE_STREAMK, -1)
    print(json.dumps(row), flush=True)
