"""tools/mg_check.py -- run under torch.distributed.run on >= 2 GPUs: checks the sharded products of
accblas_b200.mg (both transports, all op flags, real and complex) against the single-GPU library on
rank 0 and against the oracle.  Prints one line `mg_check ok ...` on rank 0; any failure raises.

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


def main():
    import accblas_b200 as ab
    from accblas_b200 import mg
    from oracle import oracle as O
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device(f"cuda:{int(os.environ.get('LOCAL_RANK', rank))}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    rng = np.random.default_rng(7)
    worst, cases = 0.0, 0
    for transport in ("peer", "p2p"):
        for (oa, ob, m, n, k) in [("N", "N", 1024, 512, 768), ("T", "N", 900, 256, 512), ("N", "T", 130, 64, 64),
                                  ("H", "H", 640, 384, 256), ("N", "N", 100, 40, 30)]:
            for dt, tdt in ((O.F64, torch.float64), (O.C64, torch.complex128), (O.F32, torch.float32)):
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
    # 2-D block sharding (mg.gemm_2d) on the grids this world size allows
    for (pr, pc) in [(1, world), (world, 1)] + ([(2, world // 2)] if world >= 4 and world % 2 == 0 else []):
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
    # gemv / dot on pre-sharded operands
    m, n = 512 * world, 384
    A = O.uniform(rng, (m, n), O.F64)
    x, x2 = O.uniform(rng, (n,), O.F64), O.uniform(rng, (m,), O.F64)
    sl = slice(rank * 512, (rank + 1) * 512)
    y = mg.gemv_sharded(1.0, "N", torch.from_numpy(A[sl]).to(dev), torch.from_numpy(x).to(dev))
    assert O.gemm_error_units(y.cpu().numpy(), O.gemv(1.0, "N", A, x), A, x, n) <= 1.0
    y = mg.gemv_sharded(1.0, "T", torch.from_numpy(A[sl]).to(dev), torch.from_numpy(x2[sl]).to(dev))
    assert O.gemm_error_units(y.cpu().numpy(), O.gemv(1.0, "T", A, x2), A, x2, m) <= 1.0
    d = mg.dot_sharded(torch.from_numpy(x2[sl]).to(dev), torch.from_numpy(x2[sl]).to(dev))
    assert abs(d.item() - O.dot(x2, x2)) <= m * 2.0 ** -53 * np.linalg.norm(x2) ** 2
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        print(f"mg_check ok: world {world}, {cases} sharded gemm cases (peer + p2p), worst {worst:.4f} units; "
              f"gemv N/T and dot sharded ok", flush=True)
    dist.destroy_process_group()
    ab.lib().accblas_peer_close_all()


if __name__ == "__main__":
    main()
