#!/usr/bin/env python
"""bench.py -- the driver's measurement contract for the BLAS hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--extras]

N = 1  workload = BASELINE.json configs[1]: Double gemm n = 8192, `A <> B` (NN); one step = one
       product of two resident 8192 x 8192 f64 matrices (2 * 8192^3 flop).  value = TFLOP/s with
       the operands already in HBM; e2e = the same product through the host-buffer C ABI
       (accblas_gemm_host: pinned host A, B -> device -> C back to host inside the timed region).
N > 1  workload = configs[4]: Double gemm n = 32768 row-sharded over the ranks (A, B start and C
       ends on rank 0: scatter A row panels, broadcast B, gather C over NCCL); strong scaling.
--impl reference   the reference's llvm-native path (one cblas_dgemm RowMajor call,
       LLVM/Native/Level3.hs:85) on the host cores through oracle/openblas_ref.py (kind "port":
       the reference is Haskell and cannot be built here; OpenBLAS is the library it calls).
One JSON line on stdout (rank 0).  At N = 1 the line also carries "extras": every other BASELINE.json config
(sgemm 8192^3 NN/NT/TN, dgemm NT/TN, zgemm / cgemm 4096^3 HN, dgemv 32768^2 N/T, ddot 2^28, dgemm 1024^3 and the
own single-GPU time of the 32768^3 product), each with its time, rate, roofline fraction against a stated
denominator, a clock sample taken during the case, an oracle check on sampled rows / entries, and the cuBLAS time
of the same call made the reference's way (baseline/cublas_ref as a subprocess: a reported number, never a
dependency of the library).  --no-extras skips them (developer loop).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))

N1 = 8192           # configs[1]
NMG = 32768         # configs[4]
METRIC = "GEMM TFLOP/s f32/f64/complex + gemv/dot HBM GB/s at 1/2/4/8 B200"
DMMA_PEAK_FALLBACK = 36.9   # TFLOP/s, baseline/ubench.cu on this pool (profiles/r01_ubench.jsonl)


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except OSError:
        return {}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is not None:
            self.proc.terminate()
        sm, mx, reasons = [], 0.0, set()
        for ts, line in self.rows:
            if ts < t0 - 0.05 or ts > t1 + 0.15:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[1]))
                mx = max(mx, float(f[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def ncu_traffic(name="r02b_dgemm_nn_8192_ncu.txt"):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    `ncu --set full` summary of the same kernel and shape (profiles/; tools/ncu_capture.sh)."""
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total, seen = 0.0, 0
    try:
        with open(os.path.join(ROOT, "profiles", name)) as f:
            for line in f:
                t = line.split()
                if len(t) >= 3 and t[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum") and t[1] in unit:
                    total += float(t[2].strip("[]',")) * unit[t[1]]
                    seen += 1
    except (OSError, ValueError):
        return None
    return round(total) if seen == 2 else None


def dmma_peak_live():
    """Issue-rate peak of the FP64 tensor pipe measured on THIS box (baseline/ubench, < 1 s)."""
    exe = os.path.join(ROOT, "baseline", "ubench")
    if os.path.exists(exe):
        try:
            out = subprocess.run([exe], capture_output=True, text=True, timeout=60).stdout
            best = 0.0
            for line in out.splitlines():
                d = json.loads(line)
                if d.get("ubench") == "dmma_m8n8k4":
                    best = max(best, d["tflops"])
            if best > 0:
                return best, "baseline/ubench dmma_m8n8k4 issue rate measured live on this GPU"
        except Exception:
            pass
    return DMMA_PEAK_FALLBACK, "baseline/ubench dmma_m8n8k4 issue rate, earlier measurement on this pool (profiles/)"


# --------------------------------------------------------------------------------------
# CPU arm: the reference's llvm-native path (cblas_dgemm RowMajor) on the host cores
# --------------------------------------------------------------------------------------
def cpu_dgemm(n, rows=None, reps=3):
    """cblas_dgemm(RowMajor, N, N), the reference's llvm-native call (LLVM/Native/Level3.hs:84-87), on all host
    cores.  rows = None: the FULL n^3 product (0.8 s at n = 8192 on 16 cores); rows = r: the first r rows of A
    against the full n x n B -- the bounded sample used when the full product would take minutes (n = 32768)."""
    import numpy as np
    from oracle import openblas_ref as R
    cores = R.host_cores()
    R.set_threads(cores)
    m = n if rows is None else rows
    rng = np.random.default_rng(0x5EED0000 + 16)
    A = rng.uniform(-1, 1, size=(m, n))
    B = rng.uniform(-1, 1, size=(n, n))
    C = np.empty((m, n))
    R.gemm(1.0, "N", A, "N", B, out=C)          # warm-up (thread pool, page faults)
    ts = []
    for _ in range(reps):
        t = time.perf_counter()
        R.gemm(1.0, "N", A, "N", B, out=C)
        ts.append(time.perf_counter() - t)
    best = min(ts)
    what = f"the full {n}^3 product" if rows is None else f"rows 0..{m - 1} of A ({m} x {n}) x B {n} x {n}"
    return {"value": round(2.0 * m * n * n / best * 1e-12, 4), "unit": "TFLOP/s", "cores": cores, "kind": "port",
            "sample": f"cblas_dgemm RowMajor NN, {what}, best of {len(ts)} ({best * 1e3:.0f} ms), "
                      f"OpenBLAS {R.config().split()[1]} via scipy, {cores} threads"}, m, (A, B, C)


def run_reference(args):
    """--impl reference: K timed steps of the CPU path on the same config/metric.  N = 1: every step is the FULL
    8192^3 cblas_dgemm.  N > 1 (32768^3, ~50 s per product on 16 cores): every step is a 512-row slice of the
    product against the full 32768^2 B -- same k and n, throughput per flop is that of the whole product."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import openblas_ref as R
    n = N1 if args.gpus == 1 else args.n_mg
    info, rows, (A, B, C) = cpu_dgemm(n, rows=None if args.gpus == 1 else min(512, n), reps=1)
    for _ in range(max(args.warmup - 1, 0)):
        R.gemm(1.0, "N", A, "N", B, out=C)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        R.gemm(1.0, "N", A, "N", B, out=C)
    dt = time.perf_counter() - t0
    tf = 2.0 * rows * n * n * args.steps / dt * 1e-12
    info["value"] = round(tf, 4)
    info["sample"] = ((f"each step = the full {n}^3 cblas_dgemm RowMajor NN" if rows == n else
                       f"each step = cblas_dgemm RowMajor NN on rows 0..{rows - 1} of A ({rows} x {n}) x B {n} x {n}: "
                       f"a bounded sample of the {n}^3 product (same k, n; {2.0 * n ** 3 / (tf * 1e12):.0f} s for the whole)")
                      + f"; OpenBLAS via scipy, {info['cores']} threads")
    line = {"impl": "reference", "metric": METRIC, "value": round(tf, 4), "unit": "TFLOP/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
            "higher_is_better": True, "scaling": "weak" if args.gpus == 1 else "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic uniform(-1,1)",
            "config": workload_config(args.gpus) | ({"m": n, "n": n, "k": n} if args.gpus > 1 else {}),
            "cpu_baseline": info,
            "e2e": {"value": round(tf, 4), "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(ngpu):
    if ngpu == 1:
        return {"workload": f"Double gemm n={N1} NN via (<>) on 1xB200 (BASELINE.json configs[1])",
                "m": N1, "n": N1, "k": N1, "opA": "N", "opB": "N", "alpha": 1,
                "l2": "inputs (1 GiB) exceed the 126 MB L2; no flush needed"}
    return {"workload": f"Double gemm n={NMG} row-sharded over {ngpu} B200 (BASELINE.json configs[4]); "
                        "A, B start and C ends on rank 0", "m": NMG, "n": NMG, "k": NMG, "opA": "N", "opB": "N",
            "alpha": 1, "parallelism": f"rows of A/C split {ngpu} ways; B broadcast (NCCL); A panels pulled by copy engines and C tiles "
                                    "stored into rank 0's HBM by the gemm epilogue over NVLink (CUDA IPC peer memory)",
            "l2": "inputs (16 GiB) exceed the 126 MB L2; no flush needed"}


# --------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------
def seeded(shape, seed, device):
    """uniform(-1,1) f64 from a fixed Philox seed."""
    import torch
    g = torch.Generator(device=device).manual_seed(seed)
    return torch.rand(shape, dtype=torch.float64, device=device, generator=g) * 2 - 1


def run_single(args):
    import numpy as np
    import torch
    import accblas_b200 as ab
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    n = N1
    flops = 2.0 * n * n * n
    A = seeded((n, n), 0x5EED0000 + 16 + 0, dev)
    B = seeded((n, n), 0x5EED0000 + 16 + 1, dev)
    assert ab.kernel_name(1, "N", "N", A.shape, B.shape) == "dgemm_dmma_tma"
    persistent = ab.workspace_bytes(1, "N", "N", A.shape, B.shape) > 0
    kernel_desc = ("dgemm_dmma_streamk_kernel<true,false> (persistent: 27 whole-tile rounds + stream-K tail, 128x128 tiles)"
                   if persistent else "dgemm_dmma_kernel<true,false,Tile128x128>")

    # ---- device-resident: W warm-ups, then exactly K timed steps ----
    for _ in range(args.warmup):
        C = ab.mXm(A, B)
    torch.cuda.synchronize()
    sampler = ClockSampler(0).start()
    time.sleep(0.25)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    l0 = ab.launch_count()
    torch.cuda.synchronize()
    t0 = time.time()
    ev[0].record()
    for i in range(args.steps):
        C = ab.mXm(A, B)
        ev[i + 1].record()
    torch.cuda.synchronize()
    t1 = time.time()
    launches = ab.launch_count() - l0
    clocks = sampler.stop(t0, t1)
    total_ms = ev[0].elapsed_time(ev[-1])
    per_step = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    kernel_ms = sum(per_step) / len(per_step)        # one kernel per step: its average launch duration
    value = flops * args.steps / total_ms * 1e-9

    # spot parity inside the bench (8 rows against the oracle): a fast wrong kernel is not a result
    from oracle import oracle as O
    rows = [0, 1, 4095, 8191]
    Ah, Bh = A[rows].cpu().numpy(), B.cpu().numpy()
    err = O.gemm_error_units(C[rows].cpu().numpy(), O.gemm(1.0, "N", Ah, "N", Bh), Ah, Bh, n)
    assert err <= 1.0, f"parity lost in bench: {err}"

    # ---- end to end through the host-buffer C ABI (pinned host memory) ----
    Ah = torch.empty((n, n), dtype=torch.float64).pin_memory()
    Bh = torch.empty((n, n), dtype=torch.float64).pin_memory()
    Ch = torch.empty((n, n), dtype=torch.float64).pin_memory()
    Ah.copy_(A)
    Bh.copy_(B)
    L = ab.lib()
    import ctypes
    one = np.array([1.0])

    def e2e_step():
        st = L.accblas_gemm_host(None, 1, 0, 0, n, n, n, n, one.ctypes.data_as(ctypes.c_void_p),
                                 ctypes.c_void_p(Ah.data_ptr()), ctypes.c_void_p(Bh.data_ptr()),
                                 ctypes.c_void_p(Ch.data_ptr()))
        assert st == 0, st
    e2e_steps = args.steps
    for _ in range(2):
        e2e_step()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()                                   # synchronous: returns with C in host memory
    e2e_s = (time.perf_counter() - t) / e2e_steps
    # the host path multiplies sub-blocks (other tile counts, other K walking order per tile): same tolerance
    # units as the parity check, a hundredth of the bound
    d_units = float(torch.linalg.norm(Ch[rows] - C[rows].cpu())) / (
        n * 2.0 ** -53 * float(torch.linalg.norm(A[rows])) * float(torch.linalg.norm(B)))
    assert d_units <= 0.01, f"e2e result differs from the device-resident result by {d_units} units"
    del Ah, Bh, Ch

    peak, peak_src = dmma_peak_live()
    cpu, _, _ = cpu_dgemm(n, reps=3)
    mp = measured_peaks()
    del A, B, C
    torch.cuda.empty_cache()
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": "TFLOP/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(total_ms / args.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic uniform(-1,1), fixed Philox seeds",
        "config": workload_config(1),
        "roofline": {"bound": "tensor", "achieved": round(flops / kernel_ms * 1e-9, 3), "peak": round(peak, 2),
                     "unit": "TFLOP/s", "frac": round(flops / kernel_ms * 1e-9 / peak, 4), "traffic": ncu_traffic(),
                     "traffic_unit": "bytes per launch (dram read + write, profiles/r02b_dgemm_nn_8192_ncu.txt); "
                                     "algorithmic bytes 1 610 612 736",
                     "kernel": kernel_desc, "flops_per_launch": flops,
                     "peak_source": peak_src + "; MEASURED_PEAKS.json carries no FP64 figure "
                                    f"(its bf16 {mp.get('bf16_tflops')} TFLOP/s is a different pipe); nominal FP64 37-40"},
        "cpu_baseline": cpu,
        "e2e": {"value": round(flops / e2e_s * 1e-12, 3), "unit": "TFLOP/s", "h2d_bytes_per_step": 2 * 8 * n * n,
                "d2h_bytes_per_step": 8 * n * n, "ms_per_step": round(e2e_s * 1e3, 3), "steps": e2e_steps,
                "api": "accblas_gemm_host (pinned host buffers)"},
        "gpu_launches": int(launches), "clocks": clocks,
        "parity": {"error_units": round(err, 5), "tolerance_units": 1.0, "rows_checked": len(rows)},
    }
    if args.extras:
        line["extras"] = extras(peak, peak_src, mp, total_ms / args.steps)
        line["summary"] = summary_of(line["extras"])
    print(json.dumps(line), flush=True)


def summary_of(ex):
    """Every BASELINE.json config of the extras in a few hundred bytes, as the LAST key of the line (a record that keeps
    only the tail of stdout still holds it): case -> [ms, rate (TFLOP/s or GB/s), roofline fraction, time of cuBLAS /
    own time (> 1: ahead), parity error in tolerance units, median SM MHz during the case]."""
    out = {"columns": ["ms", "tflops_or_gbs", "roofline_frac", "vs_cublas", "error_units", "sm_mhz"]}
    for e in ex:
        if e.get("case") == "reference_benchmark_shapes":
            out["ref_shapes_vs_cublas_sustained"] = e.get("summary")
            continue
        if "ms" not in e:
            continue
        name = e["case"].replace("gemm_", "").replace("32768x32768x32768_1gpu", "32768^3_T1")
        name = name.replace("8192x8192x8192", "8192^3").replace("4096x4096x4096", "4096^3").replace("1024x1024x1024", "1024^3")
        out[name] = [e["ms"], e.get("tflops", e.get("gbs")), (e.get("roofline") or {}).get("frac"), e.get("vs_cublas"),
                     (e.get("parity") or {}).get("error_units"), (e.get("clocks") or {}).get("sm_mhz")]
    return out


def run_multi(args):
    """N > 1: configs[4], Double gemm n = 32768, rows of A/C split over the ranks, B broadcast,
    C gathered; A, B start and C ends on rank 0.  Timed on the device, max over ranks."""
    import torch
    import torch.distributed as dist
    import accblas_b200 as ab
    from accblas_b200 import mg
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    # NCCL prints its version banner on stdout; the contract is ONE JSON line there
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    dist.init_process_group("nccl", device_id=dev)
    n = args.n_mg
    flops = 2.0 * n * n * n
    A = B = C = None
    if rank == 0:
        A = seeded((n, n), 0x5EED0000 + 64 + 0, dev)
        B = seeded((n, n), 0x5EED0000 + 64 + 1, dev)
        C = torch.empty((n, n), dtype=torch.float64, device=dev)

    # parity of the sharded paths BEFORE anything is timed (tools/mg_check.py, quick subset: C-ABI peer pipeline, plan / run,
    # SUMMA, gemv / dot exchanges, every rank): a mismatch fails the bench
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import mg_check
    mg_cases, mg_worst = mg_check.check(dev, rank, world, quick=True)

    if args.transport == "peer":
        comm = mg.get_comm(dev)
        plan = comm.gemm_plan("N", A, "N", B, shapeA=(n, n), shapeB=(n, n), dtype=torch.float64, out=C, root=0)

        def step():
            return plan.run(1.0)                 # accblas_mg_gemm_run: no handle exchange, no host synchronisation
    else:
        def step():
            return mg.gemm_rowsharded(1.0, "N", A, "N", B, shapeA=(n, n), shapeB=(n, n), dtype=torch.float64,
                                      device=dev, root=0, subpanels=args.subpanels, out=C, transport=args.transport)
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local).start() if rank == 0 else None
    l0 = ab.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    t1 = time.time()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    launches = torch.tensor([ab.launch_count() - l0], dtype=torch.int64, device=dev)
    dist.all_reduce(launches, op=dist.ReduceOp.SUM)
    total_ms = ms.item()
    # untimed extra: what ONE rank's panel product costs with its operands already local (explains the gap between
    # the step and pure compute: first-chunk exposure, per-chunk launches, the tokens, the tail)
    if args.transport == "peer":
        plan.destroy()
    rows = -(-n // world)
    rows = -(-rows // 128) * 128
    Ap = seeded((min(rows, n), n), 0x5EED0000 + 80 + rank, dev)
    Bl = B if rank == 0 else seeded((n, n), 0x5EED0000 + 65, dev)
    Cp = torch.empty((Ap.shape[0], n), dtype=torch.float64, device=dev)
    ab.gemm(1.0, "N", Ap, "N", Bl, out=Cp)
    torch.cuda.synchronize()
    dist.barrier()
    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    b0.record()
    for _ in range(2):
        ab.gemm(1.0, "N", Ap, "N", Bl, out=Cp)
    b1.record()
    torch.cuda.synchronize()
    bms = torch.tensor([b0.elapsed_time(b1) / 2], dtype=torch.float64, device=dev)
    dist.all_reduce(bms, op=dist.ReduceOp.MAX)
    del Ap, Cp
    if rank != 0:
        del Bl
    torch.cuda.empty_cache()
    l12 = mg_level12_extras(dev, rank, world) if args.n_mg >= 4096 else []
    if rank == 0:
        clocks = sampler.stop(t0, t1)
        # parity: 4 rows spread over the rank panels against the oracle
        from oracle import oracle as O
        rows = [1, n // 2 + 3, n - 1, (n // world) * (world - 1) + 7]
        Ah, Bh = A[rows].cpu().numpy(), B.cpu().numpy()
        err = O.gemm_error_units(C[rows].cpu().numpy(), O.gemm(1.0, "N", Ah, "N", Bh), Ah, Bh, n)
        assert err <= 1.0, f"parity lost in multi-GPU bench: {err}"
        peak, peak_src = DMMA_PEAK_FALLBACK, "baseline/ubench dmma_m8n8k4 issue rate on this pool (profiles/)"
        value = flops * args.steps / total_ms * 1e-9
        # what really crosses the root's links in the pull design: every peer reads ALL of B and its own panel of A out
        # of the root's HBM, and stores its panel of C into it
        egress = 8.0 * n * n * ((world - 1) + (world - 1) / world)
        ingress = 8.0 * n * n * (world - 1) / world
        # T1 of THIS product on rank 0 alone, same run, same box (after the parity rows above were read: it rewrites C)
        t1_ms = None
        if args.t1:
            try:
                ab.gemm(1.0, "N", A, "N", B, out=C)
                torch.cuda.synchronize()
                c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                reps = 3 if n >= 16384 else 10
                c0.record()
                for _ in range(reps):
                    ab.gemm(1.0, "N", A, "N", B, out=C)
                c1.record()
                torch.cuda.synchronize()
                t1_ms = c0.elapsed_time(c1) / reps
            except Exception:                     # an optional extra must not cost the line (e.g. out of memory)
                t1_ms = None
                torch.cuda.empty_cache()
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": "TFLOP/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(total_ms / args.steps, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic uniform(-1,1), fixed Philox seeds", "config": workload_config(world) | {"m": n, "n": n, "k": n},
            "roofline": {"bound": "tensor", "achieved": round(value / world, 3), "peak": peak, "unit": "TFLOP/s",
                         "frac": round(value / world / peak, 4), "traffic": None,
                         "kernel": "dgemm_dmma (persistent schedule) per rank, K-chunked; achieved = whole-job rate / ranks "
                                   "(includes every inter-GPU transfer)", "peak_source": peak_src,
                         "root_egress_bytes": egress, "root_ingress_bytes": ingress,
                         "link_note": "pull design: (N-1) |B| + (N-1)/N |A| leave the root, (N-1)/N |C| arrives"},
            "cpu_baseline": None,
            "e2e": {"value": round(value, 3), "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                    "note": "operands and result are device-resident on rank 0 by definition of configs[4]; the "
                            "timed region already contains every inter-GPU transfer"},
            "gpu_launches": int(launches.item()), "clocks": clocks,
            "parity": {"error_units": round(err, 5), "tolerance_units": 1.0, "rows_checked": len(rows)},
        }
        line["config"]["transport"] = args.transport
        line["config"]["api"] = "accblas_mg_gemm_plan / accblas_mg_gemm_run (C ABI, csrc/mg.cu)" if args.transport == "peer" else "mg.gemm_rowsharded (p2p)"
        line["panel_product_ms"] = round(bms.item(), 3)
        line["overhead_vs_panel_product_ms"] = round(total_ms / args.steps - bms.item(), 3)
        line["mg_check"] = {"cases": mg_cases, "worst_gemm_error_units": round(mg_worst, 5), "when": "before timing, every rank"}
        line["extras"] = l12
        # last keys: survive a record that keeps only the tail of stdout
        try:
            line["summary"] = {"ms_per_step": round(total_ms / args.steps, 3), "t1_ms": round(t1_ms, 3) if t1_ms else None,
                               "speedup_vs_t1": round(t1_ms / (total_ms / args.steps), 4) if t1_ms else None,
                               "t1_note": "the same n^3 product on rank 0 alone (accblas_gemm, operands resident), same run",
                               "panel_product_ms": round(bms.item(), 3), "tflops_per_gpu": round(value / world, 3),
                               "mg_check_cases": mg_cases, "error_units": round(err, 5),
                               "sm_mhz": (clocks or {}).get("sm_mhz")}
        except Exception as e:                    # a convenience key must never cost the line
            line["summary"] = {"error": repr(e)}
    dist.barrier()                                # the other ranks wait here while rank 0 measures T1
    dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    if rank == 0:
        print(json.dumps(line), flush=True)


def mg_level12_extras(dev, rank, world, iters=20):
    """configs[2] on N GPUs: dgemv 32768 x 32768 (N and T) with the rows of A pre-sharded over the ranks
    (moving 8 GiB of A would cost more than the product) and ddot n = 2^28 on pre-sharded slices.  The
    timed region holds the local kernel AND the exchange (all-gather of y slices / all-reduce of the
    partial y / all-reduce of one scalar, NCCL over NVLink); device events, max over ranks; GB/s =
    algorithmic bytes of the WHOLE product / that time.  Checked against the oracle on a few entries."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from accblas_b200 import mg
    from oracle import oracle as O
    m = n = 32768
    rows = m // world
    A = seeded((rows, n), 0x5EED0000 + 32 + 16 * rank, dev)
    x = seeded((n,), 0x5EED0000 + 33, dev)
    xs = seeded((rows,), 0x5EED0000 + 34 + 16 * rank, dev)
    nd = (1 << 28) // world
    u, v = seeded((nd,), 0x5EED0000 + 35 + 16 * rank, dev), seeded((nd,), 0x5EED0000 + 36 + 16 * rank, dev)
    out = []

    def timed(name, fn, nbytes):
        for _ in range(3):
            r = fn()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            r = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        out.append({"case": name, "ms": round(ms.item(), 4), "gbs": round(nbytes / ms.item() * 1e-6, 1),
                    "per_gpu_gbs": round(nbytes / ms.item() * 1e-6 / world, 1)})
        return r

    y = timed(f"f64gemv_N_{m}x{n}_rows/{world}", lambda: mg.gemv_sharded(1.0, "N", A, x), 8.0 * (m * n + m + n))
    pick = [0, rows // 2, rows - 1]
    ref = O.gemv(1.0, "N", A[pick].cpu().numpy(), x.cpu().numpy())
    got = y[[rank * rows + i for i in pick]].cpu().numpy()
    ok = float(np.max(np.abs(got - ref))) <= n * 2.0 ** -53 * float(torch.linalg.norm(x)) * float(torch.linalg.norm(A[pick], dim=1).max())
    yt = timed(f"f64gemv_T_{m}x{n}_rows/{world}", lambda: mg.gemv_sharded(1.0, "T", A, xs), 8.0 * (m * n + m + n))
    cols = [0, n // 3, n - 1]
    part = torch.from_numpy(O.gemv(1.0, "T", np.ascontiguousarray(A[:, cols].cpu().numpy()), xs.cpu().numpy())).to(dev)
    dist.all_reduce(part, op=dist.ReduceOp.SUM)
    bound = m * 2.0 ** -53 * float(torch.linalg.norm(A[:, cols], dim=0).max()) * float(torch.linalg.norm(xs)) * world
    ok = ok and float((yt[cols] - part).abs().max()) <= bound
    d = timed(f"f64dot_{1 << 28}_slices/{world}", lambda: mg.dot_sharded(u, v), 2.0 * 8 * (1 << 28) + 8)
    dref = torch.tensor([float(O.dot(u[:1 << 20].cpu().numpy(), v[:1 << 20].cpu().numpy()))], dtype=torch.float64, device=dev)
    dgot = mg.dot_sharded(u[:1 << 20], v[:1 << 20])
    dist.all_reduce(dref, op=dist.ReduceOp.SUM)
    tot = float((1 << 20) * world)
    ok = ok and abs(float(dgot) - float(dref)) <= tot * 2.0 ** -53 * tot / 3 and bool(torch.isfinite(d))
    flag = torch.tensor([1 if ok else 0], dtype=torch.int64, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    assert flag.item() == 1, "parity lost in the sharded gemv / dot extras"
    return out


def cublas_arm(cases, iters=8):
    """The reference's llvm-ptx path (cublas?gemm / ?gemv called its way, LLVM/PTX/Level3.hs:91-106, Level2.hs:109-124)
    for the same shapes, from baseline/cublas_ref run as a SUBPROCESS: a reported number beside ours, never linked
    into the library.  Returns {case name: median ms}."""
    exe = os.path.join(ROOT, "baseline", "cublas_ref")
    out = {}
    if not os.path.exists(exe):
        return out
    try:
        res = subprocess.run([exe, ",".join(cases), str(iters)], capture_output=True, text=True, timeout=240)
        for l in res.stdout.splitlines():
            if l.startswith("{"):
                d = json.loads(l)
                out[d["case"]] = d
    except Exception as e:      # the baseline is optional; its absence must not fail the product's bench
        out["error"] = repr(e)
    return out


def ref_shapes_arm():
    """The reference's OWN benchmark workload (bench/Bench/Accelerate.hs:44-58,92-125: gemm 100 / 250 / 500 / 1000 cubed and
    gemv / gevm 200x400 .. 2000x3000, four element types = 48 cases), this library and cuBLAS timed by the same C driver
    (baseline/sidebyside, a subprocess) on the same buffers: per-call events and the sustained time of 400 calls back to
    back.  Returns (cases, clocks over the whole arm)."""
    exe = os.path.join(ROOT, "baseline", "sidebyside")
    if not os.path.exists(exe):
        return [], None
    sampler = ClockSampler(0).start()
    t0 = time.time()
    rows = []
    try:
        res = subprocess.run([exe, "refshapes"], capture_output=True, text=True, timeout=300)
        for l in res.stdout.splitlines():
            if l.startswith("{"):
                rows.append(json.loads(l))
    except Exception as e:
        rows.append({"case": "ref_shapes_arm", "error": repr(e)})
    return rows, sampler.stop(t0, time.time())


def ref_shapes_summary(rows):
    """sustained time of cuBLAS / sustained time of this library per case (> 1: ahead), condensed."""
    r = {d["case"]: d["speedup_stream"] for d in rows if "speedup_stream" in d}
    if not r:
        return None
    import math
    behind = sorted((v, k) for k, v in r.items() if v < 1.0)
    return {"cases": len(r), "at_or_ahead_of_cublas": sum(1 for v in r.values() if v >= 1.0),
            "within_2pct": sum(1 for v in r.values() if v >= 0.98),
            "geomean": round(math.exp(sum(math.log(v) for v in r.values()) / len(r)), 4),
            "min": round(min(r.values()), 3), "max": round(max(r.values()), 3),
            "behind": {k: round(v, 3) for v, k in behind[:8]}}


def extras(dmma_peak, dmma_src, mp, headline_ms):
    """Every other BASELINE.json config at N = 1 (configs[0..3] and the T1 of configs[4]); see the module docstring."""
    import numpy as np
    import torch
    import accblas_b200 as ab
    from oracle import oracle as O
    dev = torch.device("cuda:0")
    hbm = float(mp.get("hbm_gbs") or 6553.0)
    bf16 = float(mp.get("bf16_tflops") or 1598.1)
    tf32x3_peak = bf16 / 2.0 / 3.0
    TD = {"f32": torch.float32, "f64": torch.float64, "c32": torch.complex64, "c64": torch.complex128}
    CODE = {"f32": O.F32, "f64": O.F64, "c32": O.C32, "c64": O.C64}
    ES = {"f32": 4, "f64": 8, "c32": 8, "c64": 16}
    CB = {"f32": "s", "f64": "d", "c32": "c", "c64": "z"}

    def rnd(shape, dt, seed):
        g = torch.Generator(device=dev).manual_seed(seed)
        base = torch.float32 if dt in ("f32", "c32") else torch.float64
        if dt in ("c32", "c64"):
            return torch.view_as_complex(torch.rand(tuple(shape) + (2,), dtype=base, device=dev, generator=g) * 2 - 1)
        return torch.rand(shape, dtype=base, device=dev, generator=g) * 2 - 1

    def timed(fn, min_s, est_ms, lo=3, hi=400):
        """median / best per-call device time, calls enqueued back to back (events between calls), with the
        clocks sampled during the run; enough iterations for >= min_s seconds so the 100 ms sampler sees it"""
        iters = int(max(lo, min(hi, min_s * 1e3 / max(est_ms, 1e-3))))
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        sampler = ClockSampler(0).start()
        time.sleep(0.15)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
        t0 = time.time()
        ev[0].record()
        for i in range(iters):
            fn()
            ev[i + 1].record()
        torch.cuda.synchronize()
        t1 = time.time()
        clocks = sampler.stop(t0, t1)
        ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
        return ts[len(ts) // 2], ts[0], iters, clocks

    # the headline itself, so that its cuBLAS time sits in the same record
    out = [{"case": f"f64gemm_NN_{N1}x{N1}x{N1}", "kernel": "dgemm_dmma_tma", "ms": round(headline_ms, 4),
            "tflops": round(2.0 * N1 ** 3 / headline_ms * 1e-9, 3), "note": "the headline step (value / ms_per_step of this line)",
            "cublas_case": f"dgemm_NN_{N1}x{N1}x{N1}"}]
    cublas_cases = [f"dgemm_NN_{N1}x{N1}x{N1}"]

    def gemm_case(dt, oa, ob, n, est_ms, seed, rows=4, min_s=0.45, check_cols=None):
        A, B = rnd((n, n), dt, seed), rnd((n, n), dt, seed + 1)
        name = ab.kernel_name(CODE[dt], oa, ob, A.shape, B.shape)
        box = {}
        med, best, iters, clocks = timed(lambda: box.__setitem__("C", ab.gemm(1, oa, A, ob, B)), min_s, est_ms)
        C = box["C"]
        cplx = dt[0] == "c"
        flops = (8.0 if cplx else 2.0) * n ** 3
        # parity: sampled rows (and, for the 32768^3 product, sampled columns) against the wide oracle
        ridx = [0, n - 1] + [(n // rows) * i + 17 * i for i in range(1, rows - 1)]
        it = torch.as_tensor(ridx, device=dev)
        Asub = (A[it] if oa == "N" else A[:, it].contiguous()).cpu().numpy()
        if check_cols:
            ct = torch.as_tensor(check_cols, device=dev)
            Bsub = (B[:, ct] if ob == "N" else B[ct]).contiguous().cpu().numpy()
            got = C[it][:, ct].cpu().numpy()
        else:
            Bsub = B.cpu().numpy()
            got = C[it].cpu().numpy()
        err = O.gemm_error_units(got, O.gemm(1, oa, Asub, ob, Bsub, wide=True), Asub, Bsub, n, 1)
        tol = 2.0 if ("tf32x3" in name or cplx) else 1.0
        assert err <= tol, f"parity lost in bench extras: {dt} {oa}{ob} {n}: {err}"
        rate = flops / med * 1e-9
        if dt in ("f64", "c64"):
            executed = 0.75 * rate if "3m" in name else rate      # 3M executes 6 mnk of the 8 mnk it is credited with
            roof = {"bound": "tensor", "achieved": round(executed, 3), "peak": round(dmma_peak, 2), "unit": "TFLOP/s",
                    "frac": round(executed / dmma_peak, 4), "peak_source": dmma_src,
                    "note": "executed flop rate on the FP64 tensor pipe" + (" (3M: 6mnk executed, 8mnk credited in tflops)" if "3m" in name else "")}
        else:
            executed = 0.75 * rate if "3m" in name else rate
            roof = {"bound": "tensor", "achieved": round(executed, 3), "peak": round(tf32x3_peak, 1), "unit": "TFLOP/s",
                    "frac": round(executed / tf32x3_peak, 4),
                    "peak_source": f"MEASURED_PEAKS.json bf16 {bf16} TFLOP/s / 2 (TF32 rate) / 3 (three TF32 products per FP32 product)",
                    "note": "FP32-equivalent flop rate; the tensor pipe does 3x that in TF32"}
        cname = f"{CB[dt]}gemm_{oa}{ob}_{n}x{n}x{n}"
        cublas_cases.append(cname)
        out.append({"case": f"{dt}gemm_{oa}{ob}_{n}x{n}x{n}", "kernel": name, "ms": round(med, 4), "ms_best": round(best, 4),
                    "iters": iters, "tflops": round(rate, 3), "roofline": roof, "clocks": clocks,
                    "parity": {"error_units": round(err, 5), "tolerance_units": tol, "rows_checked": len(ridx),
                               "cols_checked": len(check_cols) if check_cols else n},
                    "cublas_case": cname})
        del A, B, C, box
        torch.cuda.empty_cache()

    for ops in ("NT", "TN"):
        gemm_case("f64", ops[0], ops[1], N1, 31.0, 0x5EED0000 + 16)
    for ops in ("NN", "NT", "TN"):
        gemm_case("f32", ops[0], ops[1], N1, 4.6, 0x5EED0000 + 24)
    gemm_case("c64", "H", "N", 4096, 12.5, 0x5EED0000 + 48)
    gemm_case("c32", "H", "N", 4096, 2.3, 0x5EED0000 + 56)
    gemm_case("f64", "N", "N", 1024, 0.08, 0x5EED0000 + 0)

    # ---- configs[2]: dgemv 32768^2 N / T, ddot 2^28 (HBM roofline; denominator = MEASURED_PEAKS.json copy bandwidth)
    m = n = 32768
    A, x = rnd((m, n), "f64", 0x5EED0000 + 32), rnd((n,), "f64", 0x5EED0000 + 33)
    for op in ("N", "T"):
        box = {}
        med, best, iters, clocks = timed(lambda: box.__setitem__("y", ab.gemv(1, op, A, x)), 0.45, 1.2)
        y = box["y"]
        pick = [0, n - 1] + [(1637 * i) % n for i in range(1, 15)]
        it = torch.as_tensor(pick, device=dev)
        sub = (A[it] if op == "N" else A[:, it].contiguous()).cpu().numpy()
        ref = O.gemv(1, op, sub, x.cpu().numpy(), wide=True)
        scale = np.linalg.norm(sub, axis=1 if op == "N" else 0)
        dev_units = float(np.max(np.abs(y[it].cpu().numpy() - ref) / ((n + 2) * 2.0 ** -53 * scale * float(torch.linalg.norm(x)))))
        assert dev_units <= 1.0, f"parity lost in bench extras: dgemv {op}: {dev_units}"
        nbytes = 8.0 * (m * n + m + n)
        gbs = nbytes / med * 1e-6
        cublas_cases.append(f"dgemv_{op}_{m}x{n}")
        out.append({"case": f"f64gemv_{op}_{m}x{n}", "ms": round(med, 4), "ms_best": round(best, 4), "iters": iters,
                    "gbs": round(gbs, 1), "clocks": clocks,
                    "roofline": {"bound": "hbm", "achieved": round(gbs, 1), "peak": hbm, "unit": "GB/s", "frac": round(gbs / hbm, 4),
                                 "peak_source": "MEASURED_PEAKS.json hbm_gbs (torch copy, read + write); a read-only stream runs "
                                                "above a copy, so frac can exceed 1", "algorithmic_bytes": nbytes},
                    "parity": {"error_units": round(dev_units, 5), "tolerance_units": 1.0, "entries_checked": len(pick)},
                    "cublas_case": f"dgemv_{op}_{m}x{n}"})
    del A, x, box, y
    torch.cuda.empty_cache()
    nd = 1 << 28
    u, v = rnd((nd,), "f64", 0x5EED0000 + 35), rnd((nd,), "f64", 0x5EED0000 + 36)
    box = {}
    med, best, iters, clocks = timed(lambda: box.__setitem__("d", ab.dotu(u, v)), 0.45, 0.62)
    seg = 1 << 22                                                    # oracle on 2^22-element slices (the full vectors are checked in tests/)
    dref = float(O.dot(u[:seg].cpu().numpy(), v[:seg].cpu().numpy(), wide=True))
    dgot = float(ab.dotu(u[:seg], v[:seg]))
    du = abs(dgot - dref) / (seg * 2.0 ** -53 * float(torch.linalg.norm(u[:seg])) * float(torch.linalg.norm(v[:seg])))
    assert du <= 1.0 and bool(torch.isfinite(box["d"])), f"parity lost in bench extras: ddot {du}"
    nbytes = 16.0 * nd + 8
    gbs = nbytes / med * 1e-6
    cublas_cases.append(f"ddot_{nd}")
    out.append({"case": f"f64dot_{nd}", "ms": round(med, 4), "ms_best": round(best, 4), "iters": iters, "gbs": round(gbs, 1),
                "clocks": clocks,
                "roofline": {"bound": "hbm", "achieved": round(gbs, 1), "peak": hbm, "unit": "GB/s", "frac": round(gbs / hbm, 4),
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs (torch copy)", "algorithmic_bytes": nbytes},
                "parity": {"error_units": round(du, 6), "tolerance_units": 1.0, "entries_checked": seg,
                           "note": "first 2^22 elements against the wide oracle; the full 2^28 check is tests/test_round2_gpu.py"},
                "cublas_case": f"ddot_{nd}", "cublas_note": "stand-in: the reference's dot is Accelerate's own fold, not cuBLAS"})
    del u, v, box
    torch.cuda.empty_cache()

    # ---- T1 of configs[4]: the 32768^3 product on ONE GPU (the strong-scaling numerator for SCALE's N = 2, 4, 8)
    try:
        gemm_case("f64", "N", "N", NMG, 1950.0, 0x5EED0000 + 64, rows=4, min_s=5.0,
                  check_cols=[0, 1, 4097, 16384, 20001, 32767])
        out[-1]["case"] = "f64gemm_NN_32768x32768x32768_1gpu"
    except torch.cuda.OutOfMemoryError:
        out.append({"case": "f64gemm_NN_32768x32768x32768_1gpu", "skipped": "out of memory (24 GiB of operands)"})
        torch.cuda.empty_cache()
    ab.trim_workspace()

    # ---- the cuBLAS arm, same shapes, same box, right after (our process idles meanwhile)
    cb = cublas_arm(cublas_cases)
    for e in out:
        c = cb.get(e.get("cublas_case", ""))
        if c:
            e["cublas_ms"] = c["ms_median"]
            e["vs_cublas"] = round(c["ms_median"] / e["ms"], 4)      # > 1: this library is faster
    if "error" in cb:
        out.append({"case": "cublas_arm", "error": cb["error"]})
    # ---- the reference's own benchmark shapes, both libraries in one C driver
    rs, rs_clocks = ref_shapes_arm()
    if rs:
        out.append({"case": "reference_benchmark_shapes", "source": "bench/Bench/Accelerate.hs:44-58,92-125 via baseline/sidebyside refshapes",
                    "clocks": rs_clocks, "summary": ref_shapes_summary(rs),
                    "columns": ["case", "kernel", "accblas_us_sustained", "cublas_us_sustained", "cublas/accblas"],
                    "rows": [[d.get("case"), d.get("kernel"), round(1e3 * d.get("accblas_ms_stream", 0), 2),
                              round(1e3 * d.get("cublas_ms_stream", 0), 2), d.get("speedup_stream")] for d in rs]})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--extras", dest="extras", action="store_true", default=True)
    ap.add_argument("--no-extras", dest="extras", action="store_false")
    ap.add_argument("--no-t1", dest="t1", action="store_false", default=True,
                    help="N > 1: skip the single-GPU time of the same product (rank 0 alone, after the timed region)")
    ap.add_argument("--n-mg", type=int, default=NMG, dest="n_mg")
    ap.add_argument("--subpanels", type=int, default=4)
    ap.add_argument("--transport", default="peer", choices=["peer", "p2p"],
                    help="multi-GPU data path: IPC peer memory (gather fused into the gemm epilogue) or NCCL send/recv")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args)
    if world > 1 or args.gpus > 1:
        if world == 1:
            raise SystemExit("bench.py --gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
        return run_multi(args)
    return run_single(args)


if __name__ == "__main__":
    main()
