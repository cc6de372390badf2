"""tools/mg_check.py -- run under torch.distributed.run on >= 2 GPUs: checks the sharded products of
the C ABI's accblas_mg_* entry points (driven through accblas_b200.mg, a ctypes caller) -- all op flags, real and
complex, one-shot and plan / run -- plus the p2p transport, SUMMA and the gemv / dot exchanges, against the oracle.
bench.py --gpus N runs check(quick=True) before timing and fails on a mismatch.  Prints one line `mg_check ok ...` on rank 0; any failure raises.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/mg_check.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def check(dev, rank, world, quick=False):
    """The sharded products against the oracle: accblas_mg_gemm through the C ABI (one-shot and plan / run, every op
    class, three dtypes, K long enough for several pull chunks), the p2p transport, SUMMA (mg.gemm_2d) on the grids
    this world size allows, and the gemv / dot exchanges over peer memory.  Returns (cases, worst error units);
    raises on the first mismatch.  `quick` = the subset bench.py runs before timing."""
    import accblas_b200 as ab  # noqa: F401
    from accblas_b200 import mg
    from oracle import oracle as O
    rng = np.random.default_rng(7)
    worst, cases = 0.0, 0
    shapes = [("N", "N", 1024, 512, 768), ("T", "N", 900, 256, 512), ("N", "T", 130, 64, 64), ("H", "H", 640, 384, 256),
              ("N", "N", 100, 40, 30), ("N", "N", 640, 256, 8704), ("T", "H", 512, 200, 6000)]
    if quick:
        shapes = [("N", "N", 1024, 512, 768), ("H", "H", 640, 384, 4608), ("N", "N", 100, 40, 30)]
    for transport in ("peer",) if quick else ("peer", "p2p"):
        for (oa, ob, m, n, k) in shapes:
            for dt, tdt in ((O.F64, torch.float64), (O.C64, torch.complex128), (O.F32, torch.float32)):
                if quick and dt == O.F32:
                    continue
                A = O.uniform(rng, (m, k) if oa == "N" else (k, m), dt)
                B = O.uniform(rng, (k, n) if ob == "N" else (n, k), dt)
                tA = torch.from_numpy(A).to(dev) if rank == 0 else None
                tB = torch.from_numpy(B).to(dev) if rank == 0 else None
                for rep in range(2):          # second call exercises the cached mappings
                    C = mg.gemm_rowsharded(0.5, oa, tA, ob, tB, shapeA=A.shape, shapeB=B.shape, dtype=tdt,
                                           device=dev, transport=transport, subpanels=2)
                    torch.cuda.synchronize()
                    if rank == 0:
                        err = O.gemm_error_units(C.cpu().numpy(), O.gemm(0.5, oa, A, ob, B), A, B, k, 0.5)
                        assert err <= 2.0, (transport, oa, ob, m, n, k, dt, err)
                        worst = max(worst, err)
                        cases += 1
                    else:
                        assert C is None
    # plan / run: one plan, three products into the same C (the root rewrites A between them)
    comm = mg.get_comm(dev)
    m, n, k = 768, 640, 5120
    A = O.uniform(rng, (m, k), O.F64)
    B = O.uniform(rng, (k, n), O.F64)
    tA = torch.from_numpy(A).to(dev) if rank == 0 else None
    tB = torch.from_numpy(B).to(dev) if rank == 0 else None
    plan = comm.gemm_plan("N", tA, "N", tB, shapeA=A.shape, shapeB=B.shape, dtype=torch.float64)
    for rep in range(3):
        if rank == 0:
            tA.mul_(-1.0)
        C = plan.run(2.0)
        torch.cuda.synchronize()
        if rank == 0:
            sign = -1.0 if rep % 2 == 0 else 1.0
            err = O.gemm_error_units(C.cpu().numpy(), O.gemm(2.0 * sign, "N", A, "N", B), A, B, k, 2.0)
            assert err <= 1.0, ("plan/run", rep, err)
            cases += 1
    plan.destroy()
    # a plan the ROOT cannot make (its C is host memory: no CUDA IPC handle exists for it) fails on EVERY rank -- before
    # round 3 the root returned early and the peers waited in the handle broadcast forever -- and leaves the communicator
    # usable (the exchanges further down run on it)
    if world > 1 and not quick:
        import ctypes as ct
        bogus = np.zeros(m * n)
        hnd = ct.c_void_p()
        root_ptr = lambda t: ct.c_void_p(t.data_ptr()) if rank == 0 else ct.c_void_p(0)   # noqa: E731
        with torch.cuda.device(dev):
            st = comm.L.accblas_mg_gemm_plan(comm.handle, 1, 0, 0, m, k, k, n, root_ptr(tA), root_ptr(tB),
                                             ct.c_void_p(bogus.ctypes.data if rank == 0 else 0), 0, ct.byref(hnd))
        assert st != 0 and not hnd.value, ("plan with a host C must fail on every rank", rank, st)
        cases += 1
    # 2-D block sharding (mg.gemm_2d, SUMMA K-loop) on the grids this world size allows
    grids = [(1, world), (world, 1)] + ([(2, world // 2)] if world >= 4 and world % 2 == 0 else [])
    for (pr, pc) in grids[:1] if quick else grids:
        gi, gj = divmod(rank, pc)
        for (oa, ob, mb, nb, kb) in [("N", "N", 256, 192, 128), ("T", "N", 130, 64, 96), ("H", "H", 64, 72, 200)]:
            for dt, tdt in ((O.F64, torch.float64), (O.C32, torch.complex64)):
                m, n, k = mb * pr, nb * pc, kb * pr * pc
                A = O.uniform(rng, (m, k) if oa == "N" else (k, m), dt)
                B = O.uniform(rng, (k, n) if ob == "N" else (n, k), dt)
                opA = A if oa == "N" else A.T
                opB = B if ob == "N" else B.T
                a_blk = opA[gi * mb:(gi + 1) * mb, gj * (k // pc):(gj + 1) * (k // pc)]
                b_blk = opB[gi * (k // pr):(gi + 1) * (k // pr), gj * nb:(gj + 1) * nb]
                a_blk = np.ascontiguousarray(a_blk if oa == "N" else a_blk.T)
                b_blk = np.ascontiguousarray(b_blk if ob == "N" else b_blk.T)
                C = mg.gemm_2d(0.5, oa, torch.from_numpy(a_blk).to(dev), ob, torch.from_numpy(b_blk).to(dev), grid=(pr, pc))
                ref = O.gemm(0.5, oa, A, ob, B)[gi * mb:(gi + 1) * mb, gj * nb:(gj + 1) * nb]
                eps = 2.0 ** -53 if dt == O.F64 else 2.0 ** -24
                err = float(np.linalg.norm(C.cpu().numpy() - ref)) / ((k + 2) * eps * 0.5 * np.linalg.norm(A) * np.linalg.norm(B))
                assert err <= 2.0, ("gemm_2d", pr, pc, oa, ob, dt, err)
                cases += 1
    # gemv / dot on pre-sharded operands: exchange over peer memory; every rank must hold the same bits
    for dt, tdt in ((O.F64, torch.float64), (O.C32, torch.complex64)):
        m, n = 512 * world, 384
        A = O.uniform(rng, (m, n), dt)
        x, x2 = O.uniform(rng, (n,), dt), O.uniform(rng, (m,), dt)
        sl = slice(rank * 512, (rank + 1) * 512)
        for rep in range(3):      # back-to-back collectives alternate the two halves of the exchange buffer
            y = mg.gemv_sharded(1.0, "N", torch.from_numpy(A[sl]).to(dev), torch.from_numpy(x).to(dev))
            assert O.gemm_error_units(y.cpu().numpy(), O.gemv(1.0, "N", A, x), A, x, n) <= 2.0
            # panels as accblas_mg_row_panel cuts them: unequal (the last one short) and, for m = 256, EMPTY on ranks >= 2
            for mt in (m - 24, 256):
                s0, r0 = mg.row_panels(mt, world)[rank]
                y = mg.gemv_sharded(1.0, "N", torch.from_numpy(A[s0:s0 + r0]).to(dev), torch.from_numpy(x).to(dev), m_total=mt)
                assert y.shape == (mt,)
                assert O.gemm_error_units(y.cpu().numpy(), O.gemv(1.0, "N", A[:mt], x), A[:mt], x, n) <= 2.0
            for op in ("T", "H"):
                y = mg.gemv_sharded(1.0, op, torch.from_numpy(A[sl]).to(dev), torch.from_numpy(x2[sl]).to(dev))
                assert O.gemm_error_units(y.cpu().numpy(), O.gemv(1.0, op, A, x2), A, x2, m) <= 2.0
            for conj in (False, True):
                d = mg.dot_sharded(torch.from_numpy(x2[sl]).to(dev), torch.from_numpy(x2[sl]).to(dev), conj=conj)
                ref = O.dot(x2, x2, conj=conj)
                assert abs(complex(d.item()) - complex(ref)) <= m * O.EPS[dt] * np.linalg.norm(x2) ** 2
        same = torch.view_as_real(y).clone() if y.is_complex() else y.clone()
        lo, hi = same.clone(), same.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert torch.equal(lo, hi), "ranks hold different bits of the all-reduced y"
        cases += 1
    torch.cuda.synchronize()
    dist.barrier()
    return cases, worst


def main():
    import accblas_b200 as ab
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device(f"cuda:{int(os.environ.get('LOCAL_RANK', rank))}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    cases, worst = check(dev, rank, world)
    if rank == 0:
        print(f"mg_check ok: world {world}, {cases} sharded cases (C-ABI peer pipeline + p2p + SUMMA + gemv / dot exchange), "
              f"worst gemm error {worst:.4f} units", flush=True)
    dist.destroy_process_group()
    ab.lib().accblas_peer_close_all()


if __name__ == "__main__":
    main()
