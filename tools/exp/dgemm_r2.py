"""tools/exp/dgemm_r2.py -- round-2 dgemm experiment: stage-release mode x schedule (data-parallel / whole-tile rounds +
stream-K tail / pure stream-K) on the BASELINE shape and the shapes where cuBLAS led in round 1; then the small-tile
race shapes under the light release (does __syncwarp alone hold where the proxy fence was introduced?)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, ROOT)
import perf  # noqa: E402
import torch  # noqa: E402
import accblas_b200 as ab  # noqa: E402

SCHED = {"dp": 0, "hybrid": 1, "streamk": 2}
for (n, ops, iters) in [(8192, "NN", 8), (8192, "NT", 5), (8192, "TN", 5), (8192, "TT", 5), (2048, "NN", 30)]:
    A, B = perf.rnd((n, n), "f64", 1), perf.rnd((n, n), "f64", 2)
    row = {"sweep": f"f64gemm_{ops}", "n": n}
    for rel in (0, 1, 2):
        for name, v in SCHED.items():
            ab.set_tuning(ab.TUNE_RELEASE, rel)
            ab.set_tuning(ab.TUNE_STREAMK, v)
            med, best = perf.timeit(lambda: ab.gemm(1, ops[0], A, ops[1], B), iters)
            row[f"{name}_rel{rel}_ms"] = round(med, 4)
    if ops != "NT":                         # MN-major operand panels as 16 x 16 boxes (round 1) instead of one 3-D box
        ab.set_tuning(ab.TUNE_MN_ONEBOX, 0)
        for rel in (0, 2):
            ab.set_tuning(ab.TUNE_RELEASE, rel)
            for name, v in (("dp", 0), ("hybrid", 1)):
                ab.set_tuning(ab.TUNE_STREAMK, v)
                med, best = perf.timeit(lambda: ab.gemm(1, ops[0], A, ops[1], B), iters)
                row[f"{name}_rel{rel}_boxes16_ms"] = round(med, 4)
        ab.set_tuning(ab.TUNE_MN_ONEBOX, -1)
    ab.set_tuning(ab.TUNE_RELEASE, -1)
    ab.set_tuning(ab.TUNE_STREAMK, -1)
    med, best = perf.timeit(lambda: ab.gemm(1, ops[0], A, ops[1], B), iters)
    row["auto_ms"] = round(med, 4)
    row["auto_tflops"] = round(2.0 * n ** 3 / med * 1e-9, 3)
    print(json.dumps(row), flush=True)
    del A, B

# zgemm 4096^3 HN: 4M big tile under both releases, and 3M
A, B = perf.rnd((4096, 4096), "c64", 1), perf.rnd((4096, 4096), "c64", 2)
row = {"sweep": "c64gemm_HN", "n": 4096}
for scheme in (0, 1):
    for rel in (0, 2):
        ab.set_tuning(ab.TUNE_ZGEMM_3M, scheme)
        ab.set_tuning(ab.TUNE_RELEASE, rel)
        med, best = perf.timeit(lambda: ab.gemm(1, "H", A, "N", B), 5)
        row[f"{'3m' if scheme else '4m'}_rel{rel}_ms"] = round(med, 4)
ab.set_tuning(ab.TUNE_ZGEMM_3M, -1)
ab.set_tuning(ab.TUNE_RELEASE, -1)
print(json.dumps(row), flush=True)
del A, B

# the race shapes (tests/test_parity_gpu.py::test_gemm_many_small_tiles_short_k) under the LIGHT release on the small tiles
from oracle import oracle as O  # noqa: E402
import numpy as np  # noqa: E402
for dt, tdt, knob, tile in ((O.F64, torch.float64, ab.TUNE_DMMA_TILE, 3), (O.C64, torch.complex128, ab.TUNE_ZMMA_TILE, 2)):
    bad = {0: 0, 2: 0}
    runs = 0
    ab.set_tuning(knob, tile)
    ab.set_tuning(ab.TUNE_ZGEMM_3M, 0)
    for (m, n, k) in [(768, 768, 96), (1024, 1024, 64), (2048, 2048, 48)]:
        for oa, ob in [("N", "N"), ("T", "N"), ("N", "H")]:
            A = perf.rnd((m, k) if oa == "N" else (k, m), "f64" if dt == O.F64 else "c64", 7)
            B = perf.rnd((k, n) if ob == "N" else (n, k), "f64" if dt == O.F64 else "c64", 8)
            ab.set_tuning(ab.TUNE_RELEASE, 0)
            ref = ab.gemm(1, oa, A, ob, B)
            for rel in (0, 2):
                ab.set_tuning(ab.TUNE_RELEASE, rel)
                for rep in range(40):
                    C = ab.gemm(1, oa, A, ob, B)
                    if not torch.equal(C, ref):
                        bad[rel] += 1
                    runs += 1
    ab.set_tuning(knob, -1)
    ab.set_tuning(ab.TUNE_ZGEMM_3M, -1)
    ab.set_tuning(ab.TUNE_RELEASE, -1)
    print(json.dumps({"race_check": O.NAMES[dt], "small_tile": tile, "runs": runs, "mismatches_fenced": bad[0],
                      "mismatches_delayed": bad[2]}), flush=True)
