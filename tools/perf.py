"""tools/perf.py -- developer timing loop (CUDA events on torch's current stream, which is the
stream the C ABI launches on).  Prints one JSON line per case.
usage: python tools/perf.py gemm f64 N N 8192 8192 8192 [iters]
       python tools/perf.py gemv f64 N 32768 32768 [iters]
       python tools/perf.py dot f64 268435456 [iters]
       python tools/perf.py suite
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
import torch  # noqa: E402
import accblas_b200 as ab  # noqa: E402

TD = {"f32": torch.float32, "f64": torch.float64, "c32": torch.complex64, "c64": torch.complex128}
CODE = {"f32": 0, "f64": 1, "c32": 2, "c64": 3}
ES = {"f32": 4, "f64": 8, "c32": 8, "c64": 16}


def rnd(shape, dt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    base = torch.float32 if dt in ("f32", "c32") else torch.float64
    if dt in ("c32", "c64"):
        v = torch.rand(tuple(shape) + (2,), dtype=base, device="cuda", generator=g) * 2 - 1
        return torch.view_as_complex(v)
    return torch.rand(shape, dtype=base, device="cuda", generator=g) * 2 - 1


def timeit(fn, iters, warm=3):
    """Per-call device time with the calls enqueued back to back (events between calls, one
    synchronize at the end): the GPU never idles waiting for Python, so a launch-latency gap is not
    billed to the kernel.  When the host is the slower side (tiny products) this is the API's
    sustained time per call."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return ts[len(ts) // 2], ts[0]


def gemm_case(dt, oa, ob, m, n, k, iters=10):
    A = rnd((m, k) if oa == "N" else (k, m), dt, 1)
    B = rnd((k, n) if ob == "N" else (n, k), dt, 2)
    med, best = timeit(lambda: ab.gemm(1, oa, A, ob, B), iters)
    fl = (8.0 if dt[0] == "c" else 2.0) * m * n * k
    print(json.dumps({"case": f"{dt}gemm_{oa}{ob}_{m}x{n}x{k}", "kernel": ab.kernel_name(CODE[dt], oa, ob, A.shape, B.shape),
                      "ms_median": round(med, 4), "ms_best": round(best, 4),
                      "tflops_median": round(fl / med * 1e-9, 3), "tflops_best": round(fl / best * 1e-9, 3)}), flush=True)


def gemv_case(dt, oa, m, n, iters=30):
    A = rnd((m, n), dt, 3)
    x = rnd((n if oa == "N" else m,), dt, 4)
    med, best = timeit(lambda: ab.gemv(1, oa, A, x), iters)
    by = ES[dt] * (m * n + m + n)
    print(json.dumps({"case": f"{dt}gemv_{oa}_{m}x{n}", "ms_median": round(med, 4), "ms_best": round(best, 4),
                      "gbs_median": round(by / med * 1e-6, 1), "gbs_best": round(by / best * 1e-6, 1)}), flush=True)


def dot_case(dt, n, iters=30):
    x, y = rnd((n,), dt, 5), rnd((n,), dt, 6)
    med, best = timeit(lambda: ab.dotu(x, y), iters)
    by = 2 * ES[dt] * n + ES[dt]
    print(json.dumps({"case": f"{dt}dot_{n}", "ms_median": round(med, 4), "ms_best": round(best, 4),
                      "gbs_median": round(by / med * 1e-6, 1), "gbs_best": round(by / best * 1e-6, 1)}), flush=True)


def l1_case(name, dt, n, iters=30):
    x, y = rnd((n,), dt, 5), rnd((n,), dt, 6)
    fn = {"asum": lambda: ab.asum(x), "amax": lambda: ab.amax(x), "amin": lambda: ab.amin(x),
          "sdot": lambda: ab.sdot(0.5, x, y)}[name]
    med, best = timeit(fn, iters)
    by = (2 if name == "sdot" else 1) * ES[dt] * n
    print(json.dumps({"case": f"{dt}{name}_{n}", "ms_median": round(med, 4), "ms_best": round(best, 4),
                      "gbs_median": round(by / med * 1e-6, 1), "gbs_best": round(by / best * 1e-6, 1)}), flush=True)


def suite():
    for n in (100, 250, 500, 512, 1000, 1024, 2048, 4096):
        gemm_case("f64", "N", "N", n, n, n, 50)
    for n in (100, 250, 500, 1000):
        gemm_case("f32", "N", "N", n, n, n, 50)
    for ops in ("NN", "NT", "TN", "TT"):
        gemm_case("f64", ops[0], ops[1], 8192, 8192, 8192, 5)
    for ops in ("NN", "NT", "TN"):
        gemm_case("f32", ops[0], ops[1], 8192, 8192, 8192, 5)
    gemm_case("c64", "N", "N", 4096, 4096, 4096, 5)
    gemm_case("c64", "H", "N", 4096, 4096, 4096, 5)
    gemm_case("c32", "N", "N", 4096, 4096, 4096, 5)
    gemm_case("c32", "H", "N", 4096, 4096, 4096, 5)
    for dt in ("f64", "f32", "c64", "c32"):
        for oa in "NT":
            gemv_case(dt, oa, 32768 if dt == "f64" else 16384, 32768 if dt == "f64" else 16384)
    gemv_case("f64", "N", 2000, 3000)
    gemv_case("f64", "T", 2000, 3000)
    for dt in ("f64", "f32", "c64", "c32"):
        dot_case(dt, 1 << 28 if dt == "f64" else 1 << 27)
    dot_case("f64", 1 << 20)
    for name in ("sdot", "asum", "amax", "amin"):
        for dt in ("f64", "f32", "c64", "c32"):
            l1_case(name, dt, 1 << 28 if dt == "f64" else 1 << 27)


if __name__ == "__main__":
    a = sys.argv[1:]
    # kernel variants are forced with the library's own environment variable, e.g. ACCBLAS_TUNE=7=0,0=1
    if not a or a[0] == "suite":
        suite()
    elif a[0] == "gemm":
        gemm_case(a[1], a[2], a[3], int(a[4]), int(a[5]), int(a[6]), int(a[7]) if len(a) > 7 else 10)
    elif a[0] == "gemv":
        gemv_case(a[1], a[2], int(a[3]), int(a[4]), int(a[5]) if len(a) > 5 else 30)
    elif a[0] in ("asum", "amax", "amin", "sdot"):
        l1_case(a[0], a[1], int(a[2]), int(a[3]) if len(a) > 3 else 30)
    elif a[0] == "dot":
        dot_case(a[1], int(a[2]), int(a[3]) if len(a) > 3 else 30)
