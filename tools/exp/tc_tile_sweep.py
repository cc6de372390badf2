"""tools/exp/tc_tile_sweep.py -- f32 / c32 gemm: tcgen05 CTA tiles (ACCBLAS_TUNE_TC_TILE 0/1/2) against the
mma.sync family and the automatic choice, one JSON line per size (profiles/r01_tc_tiles.jsonl)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import perf  # noqa: E402
import accblas_b200 as ab  # noqa: E402


def t_us(dt, n, knobs):
    for k, v in knobs.items():
        ab.set_tuning(k, v)
    try:
        A, B = perf.rnd((n, n), dt, 1), perf.rnd((n, n), dt, 2)
        name = ab.kernel_name(perf.CODE[dt], "N", "N", A.shape, B.shape)
        med, _ = perf.timeit(lambda: ab.gemm(1, "N", A, "N", B), 30 if n <= 2048 else 10)
        return round(med * 1e3, 1), name
    finally:
        for k in knobs:
            ab.set_tuning(k, -1)


for dt in ("f32", "c32"):
    for n in (250, 384, 500, 768, 1000, 1024, 1536, 2048, 3072, 4096, 8192):
        if dt == "c32" and n > 4096:
            continue
        row = {"sweep": dt + "gemm_NN", "n": n}
        for tile in (0, 1, 2):
            us, name = t_us(dt, n, {2: 2, 8: tile})
            row[f"tc_tile{tile}_us"] = us if "tcgen05" in name else None
        if n <= 2048:
            us, name = t_us(dt, n, {2: 1})
            row["mma_us"] = us if "mma" in name else None
        row["auto_us"], row["auto_kernel"] = t_us(dt, n, {})
        print(json.dumps(row), flush=True)
