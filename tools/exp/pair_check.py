"""tools/exp/pair_check.py -- the CTA-pair (cta_group::2) tcgen05 tile against the one-CTA tile: correctness on ragged and
planed shapes, then timing at 4096^3 / 8192^3 (f32 NN / TN, c32 HN on the 4M and 3M paths)."""
import json
import sys
import os

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
import torch  # noqa: E402
import accblas_b200 as ab  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(1)


def rnd(shape, cx):
    a = torch.rand(shape, device=dev, dtype=torch.float32) * 2 - 1
    if cx:
        a = torch.complex(a, torch.rand(shape, device=dev, dtype=torch.float32) * 2 - 1)
    return a


def op(o, X):
    return X if o == "N" else (X.T if o == "T" else X.T.conj())


def check(m, n, k, oa, ob, cx, tile, c3m=-1):
    A = rnd((m, k) if oa == "N" else (k, m), cx)
    B = rnd((k, n) if ob == "N" else (n, k), cx)
    ab.set_tuning(ab.TUNE_TC_TILE, tile)
    ab.set_tuning(ab.TUNE_SGEMM_FAMILY, 2)
    if cx:
        ab.set_tuning(ab.TUNE_CGEMM_3M, c3m)
    try:
        C = ab.gemm(1, oa, A, ob, B)
    finally:
        ab.set_tuning(ab.TUNE_TC_TILE, -1); ab.set_tuning(ab.TUNE_SGEMM_FAMILY, -1); ab.set_tuning(ab.TUNE_CGEMM_3M, -1)
    wide = torch.complex128 if cx else torch.float64
    ref = op(oa, A).to(wide) @ op(ob, B).to(wide)
    err = ((C.to(wide) - ref).abs().max() / ref.abs().max()).item()
    return err


rows = []
for (m, n, k) in [(256, 256, 64), (512, 768, 96), (384, 256, 200), (1000, 900, 333), (257, 300, 40), (2048, 2048, 2048), (1100, 1000, 520)]:
    for (oa, ob) in [("N", "N"), ("T", "N"), ("N", "T")]:
        e3 = check(m, n, k, oa, ob, False, 3)
        rows.append({"f32": (m, n, k), "ops": oa + ob, "err_pair": e3})
        assert e3 < 2e-6, rows[-1]
for (m, n, k) in [(512, 384, 128), (1100, 1000, 520), (300, 520, 77)]:
    for c3m in (0, 1):
        e3 = check(m, n, k, "H", "N", True, 3, c3m)
        rows.append({"c32": (m, n, k), "c3m": c3m, "err_pair": e3})
        assert e3 < 4e-6, rows[-1]
print(json.dumps({"correct": len(rows), "worst": max(r["err_pair"] for r in rows)}), flush=True)


def bench(m, n, k, oa, ob, cx, tile, c3m=-1, iters=20):
    A = rnd((m, k) if oa == "N" else (k, m), cx)
    B = rnd((k, n) if ob == "N" else (n, k), cx)
    ab.set_tuning(ab.TUNE_TC_TILE, tile)
    if cx:
        ab.set_tuning(ab.TUNE_CGEMM_3M, c3m)
    out = torch.empty((m, n), device=dev, dtype=A.dtype)
    for _ in range(3):
        ab.gemm(1, oa, A, ob, B, out=out)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        ab.gemm(1, oa, A, ob, B, out=out)
        ev[i + 1].record()
    torch.cuda.synchronize()
    ab.set_tuning(ab.TUNE_TC_TILE, -1); ab.set_tuning(ab.TUNE_CGEMM_3M, -1)
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return round(ts[len(ts) // 2], 4), round(ts[0], 4)


for (n, cx, oa, ob, c3m) in [(4096, False, "N", "N", -1), (8192, False, "N", "N", -1), (8192, False, "T", "N", -1), (8192, False, "N", "T", -1),
                             (4096, True, "H", "N", 0), (4096, True, "H", "N", 1), (2048, False, "N", "N", -1)]:
    r = {"n": n, "cx": cx, "ops": oa + ob, "c3m": c3m}
    for tile, name in ((0, "one_cta"), (3, "pair"), (-1, "auto")):
        r[name] = bench(n, n, n, oa, ob, cx, tile, c3m, iters=20 if n >= 8192 else 40)
    fl = (8 if cx else 2) * n ** 3
    r["tflops_pair_median"] = round(fl / r["pair"][0] * 1e-9, 1)
    r["tflops_one_median"] = round(fl / r["one_cta"][0] * 1e-9, 1)
    print(json.dumps(r), flush=True)
