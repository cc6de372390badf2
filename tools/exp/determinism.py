"""tools/exp/determinism.py -- which (schedule, release) combination of the f64 kernel is not bit-reproducible?"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import perf  # noqa: E402
import torch  # noqa: E402
import accblas_b200 as ab  # noqa: E402

for (n, ops) in [(2048, "NN"), (2048, "TN"), (2048, "NT"), (1536, "NN"), (4096, "NN")]:
    A, B = perf.rnd((n, n), "f64", 1), perf.rnd((n, n), "f64", 2)
    ab.set_tuning(ab.TUNE_STREAMK, 0)
    ab.set_tuning(ab.TUNE_RELEASE, 0)
    base = ab.gemm(1, ops[0], A, ops[1], B)
    row = {"determinism": f"f64gemm_{ops}_{n}"}
    for sched in (0, 1, 2):
        for rel in (0, 1, 2):
            ab.set_tuning(ab.TUNE_STREAMK, sched)
            ab.set_tuning(ab.TUNE_RELEASE, rel)
            first = ab.gemm(1, ops[0], A, ops[1], B)
            diff = 0
            worst = 0.0
            for rep in range(25):
                C = ab.gemm(1, ops[0], A, ops[1], B)
                if not torch.equal(C, first):
                    diff += 1
                    worst = max(worst, float((C - first).abs().max()))
            row[f"sched{sched}_rel{rel}"] = [diff, worst, float((first - base).abs().max())]
    print(json.dumps(row), flush=True)
ab.set_tuning(ab.TUNE_STREAMK, -1)
ab.set_tuning(ab.TUNE_RELEASE, -1)
