"""tools/err_table.py -- componentwise and normwise error of the split-precision paths against a float64 product, next to
what cuBLAS's true-FP32 routines (cublasSgemm / cublasCgemm through torch.matmul with TF32 off -- a REPORTED baseline,
never part of the product) give on the same inputs.  Answers "is 3xTF32 / 3M FP32-class?" with numbers
(Parity): one JSON line per case into profiles/r02_error_table.jsonl.

  max_rel   max_ij |C - C64| / (|A| |B|)_ij          componentwise, relative to the condition-free bound
  norm_rel  ||C - C64||_F / (||A||_F ||B||_F)         normwise
both divided by the unit roundoff of float32 (2^-24): the number of "half ulps" of error per output."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
import torch  # noqa: E402
import accblas_b200 as ab  # noqa: E402

torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")
U = 2.0 ** -24


def rnd(shape, cx, gen):
    a = torch.rand(shape, device=dev, dtype=torch.float32, generator=gen) * 2 - 1
    if cx:
        a = torch.complex(a, torch.rand(shape, device=dev, dtype=torch.float32, generator=gen) * 2 - 1)
    return a


def op(o, X):
    return X if o == "N" else (X.T if o == "T" else X.T.conj())


def errors(C, A, B, oa, ob):
    wide = torch.complex128 if C.is_complex() else torch.float64
    Aw, Bw = op(oa, A).to(wide), op(ob, B).to(wide)
    ref = Aw @ Bw
    bound = Aw.abs() @ Bw.abs()
    d = (C.to(wide) - ref).abs()
    return {"max_rel_u": round((d / bound).max().item() / U, 3),
            "norm_rel_u": round((d.norm() / (Aw.norm() * Bw.norm())).item() / U, 4)}


def ours(A, B, oa, ob, tile, kcb_env, c3m):
    ab.set_tuning(ab.TUNE_TC_TILE, tile)
    if A.is_complex():
        ab.set_tuning(ab.TUNE_CGEMM_3M, c3m)
    try:
        return ab.gemm(1, oa, A, ob, B)
    finally:
        ab.set_tuning(ab.TUNE_TC_TILE, -1); ab.set_tuning(ab.TUNE_CGEMM_3M, -1)


gen = torch.Generator(device=dev); gen.manual_seed(20260)
out = []
cases = [(1024, False, "N", "N"), (4096, False, "N", "N"), (8192, False, "N", "N"), (8192, False, "T", "N"),
         (1024, True, "H", "N"), (4096, True, "H", "N")]
for (n, cx, oa, ob) in cases:
    A, B = rnd((n, n), cx, gen), rnd((n, n), cx, gen)
    row = {"case": f"{'c32' if cx else 'f32'}gemm_{oa}{ob}_{n}", "kcb": os.environ.get("ACCBLAS_TC_KCB", "default"),
           "kernel": ab.kernel_name(2 if cx else 0, oa, ob, A.shape, B.shape)}
    row["cublas_fp32"] = errors(op(oa, A) @ op(ob, B), A, B, oa, ob)
    if cx:
        row["ours_4m"] = errors(ours(A, B, oa, ob, -1, None, 0), A, B, oa, ob)
        row["ours_3m"] = errors(ours(A, B, oa, ob, -1, None, 1), A, B, oa, ob)
    else:
        row["ours"] = errors(ours(A, B, oa, ob, -1, None, -1), A, B, oa, ob)
        row["ours_one_cta_tile"] = errors(ours(A, B, oa, ob, 0, None, -1), A, B, oa, ob)
    print(json.dumps(row), flush=True)
    del A, B
# the f64 3M path against cublasZgemm (4M)
for n in (1024, 4096):
    A = torch.complex(torch.rand((n, n), device=dev, dtype=torch.float64, generator=gen) * 2 - 1, torch.rand((n, n), device=dev, dtype=torch.float64, generator=gen) * 2 - 1)
    B = torch.complex(torch.rand((n, n), device=dev, dtype=torch.float64, generator=gen) * 2 - 1, torch.rand((n, n), device=dev, dtype=torch.float64, generator=gen) * 2 - 1)
    # float64 reference: the same product in "double-double by parts" is not available; compare both to each other and
    # to the componentwise bound instead (units of 2^-53)
    U64 = 2.0 ** -53
    bound = A.abs() @ B.abs()
    ref = A.conj().T @ B                           # cublasZgemm, 4M
    row = {"case": f"c64gemm_HN_{n}", "unit": "2^-53"}
    for name, knob in (("ours_3m", 1), ("ours_4m", 0)):
        ab.set_tuning(ab.TUNE_ZGEMM_3M, knob)
        C = ab.gemm(1, "H", A, "N", B)
        ab.set_tuning(ab.TUNE_ZGEMM_3M, -1)
        row[name + "_vs_cublas_4m"] = {"max_rel_u": round(((C - ref).abs() / bound).max().item() / U64, 3)}
    print(json.dumps(row), flush=True)
