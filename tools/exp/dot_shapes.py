"""tools/exp/dot_shapes.py -- the dot streaming pass under every ACCBLAS_TUNE_DOT_SHAPE, plus the compensated dot, on
the BASELINE size (ddot 2^28) and the reference's benchmark sizes; one JSON line per (case, shape)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch  # noqa: E402
import accblas_b200 as ab  # noqa: E402
from perf import rnd, timeit, ES  # noqa: E402

SHAPES = {0: "512x4x2", 1: "256x4x4", 2: "256x8x2", 4: "512x8x1", 5: "1024x4x1"}
for dt, n, iters in (("f64", 1 << 28, 200), ("f32", 1 << 28, 200), ("c64", 1 << 27, 200), ("f64", 1 << 20, 400), ("f32", 1 << 20, 400),
                     ("c32", 1 << 20, 400)):
    x, y = rnd((n,), dt, 5), rnd((n,), dt, 6)
    for sh, name in SHAPES.items():
        ab.set_tuning(ab.TUNE_DOT_SHAPE, sh)
        med, best = timeit(lambda: ab.dotu(x, y), iters)
        print(json.dumps({"case": f"{dt}dot_{n}", "shape": name, "ms_median": round(med, 4), "ms_best": round(best, 4),
                          "gbs_median": round(2 * ES[dt] * n / med * 1e-6, 1)}), flush=True)
    ab.set_tuning(ab.TUNE_DOT_SHAPE, -1)
    med, best = timeit(lambda: ab.dot_compensated(x, y), iters)
    print(json.dumps({"case": f"{dt}dot_{n}", "shape": "compensated 512x4x2", "ms_median": round(med, 4), "ms_best": round(best, 4),
                      "gbs_median": round(2 * ES[dt] * n / med * 1e-6, 1)}), flush=True)
    del x, y
