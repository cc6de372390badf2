"""CPU, world_size 2, gloo: the sharding + collective logic of accblas_b200.mg with the local product
injected (the oracle's OpenBLAS pin stands in for the CUDA library, which needs a GPU)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _local_gemm(alpha, opa, A, opb, B, out=None, beta=None, Cin=None):
    from oracle import openblas_ref as R
    C = torch.from_numpy(R.gemm(alpha, opa, A.numpy(), opb, B.numpy()))
    if Cin is not None:
        C = C + (1 if beta is None else beta) * Cin
    if out is not None:
        out.copy_(C)
        return out
    return C


def _local_gemv(alpha, opa, A, x):
    from oracle import openblas_ref as R
    return torch.from_numpy(R.gemv(alpha, opa, A.numpy(), x.numpy()))


def _local_dot(x, y):
    from oracle import openblas_ref as R
    return torch.tensor(R.dot(x.numpy(), y.numpy()))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from accblas_b200 import mg
        from oracle import oracle as O
        rng = np.random.default_rng(42)          # same stream on both ranks: both can form the expectation
        res = {}
        for (oa, ob, m, n, k) in [("N", "N", 300, 70, 90), ("T", "N", 257, 33, 64), ("N", "T", 5, 40, 20),
                                  ("H", "H", 384, 16, 48)]:
            for dt in (O.F64, O.C64):
                A = O.uniform(rng, (m, k) if oa == "N" else (k, m), dt)
                B = O.uniform(rng, (k, n) if ob == "N" else (n, k), dt)
                tA, tB = torch.from_numpy(A), torch.from_numpy(B)
                C = mg.gemm_rowsharded(0.5, oa, tA if rank == 0 else None, ob, tB if rank == 0 else None,
                                       shapeA=A.shape, shapeB=B.shape, dtype=tA.dtype, device="cpu",
                                       subpanels=2, local_gemm=_local_gemm)
                if rank == 0:
                    ref = O.gemm(0.5, oa, A, ob, B)
                    res[f"gemm_{oa}{ob}_{dt}"] = O.gemm_error_units(C.numpy(), ref, A, B, k, 0.5)
                else:
                    assert C is None
        # gemv N: rows pre-sharded, x replicated; T: rows and x sharded
        A = O.uniform(rng, (256, 100), O.F64)
        x = O.uniform(rng, (100,), O.F64)
        half = slice(rank * 128, (rank + 1) * 128)
        y = mg.gemv_sharded(1.0, "N", torch.from_numpy(A[half]), torch.from_numpy(x), local_gemv=_local_gemv)
        res["gemv_N"] = O.gemm_error_units(y.numpy(), O.gemv(1.0, "N", A, x), A, x, 100)
        # ragged panels, as accblas_mg_row_panel cuts them: 200 rows on 2 ranks = 128 + 72, 100 rows = 100 + 0
        for mt in (200, 100):
            s0, r0 = mg.row_panels(mt, 2)[rank]
            y = mg.gemv_sharded(1.0, "N", torch.from_numpy(A[s0:s0 + r0]), torch.from_numpy(x), local_gemv=_local_gemv,
                                m_total=mt)
            assert y.shape == (mt,)
            res[f"gemv_N_ragged_{mt}"] = O.gemm_error_units(y.numpy(), O.gemv(1.0, "N", A[:mt], x), A[:mt], x, 100)
        x2 = O.uniform(rng, (256,), O.F64)
        y = mg.gemv_sharded(1.0, "T", torch.from_numpy(A[half]), torch.from_numpy(x2[half]), local_gemv=_local_gemv)
        res["gemv_T"] = O.gemm_error_units(y.numpy(), O.gemv(1.0, "T", A, x2), A, x2, 256)
        a, b = O.uniform(rng, (1000,), O.F64), O.uniform(rng, (1000,), O.F64)
        h = slice(rank * 500, (rank + 1) * 500)
        d = mg.dot_sharded(torch.from_numpy(a[h]), torch.from_numpy(b[h]), local_dot=_local_dot)
        res["dot"] = abs(d.item() - O.dot(a, b)) / (1000 * 2.0 ** -53 * np.linalg.norm(a) * np.linalg.norm(b))
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def _worker_2d(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from accblas_b200 import mg
        from oracle import oracle as O
        rng = np.random.default_rng(43)          # same stream on every rank
        res = {}
        for (pr, pc) in [(2, 2), (1, 4), (4, 1)]:
            i, j = divmod(rank, pc)
            for (oa, ob, m, n, k) in [("N", "N", 48, 40, 64), ("T", "N", 16, 24, 32), ("N", "H", 8, 12, 16), ("H", "T", 20, 4, 8)]:
                for dt in (O.F64, O.C64):
                    A = O.uniform(rng, (m, k) if oa == "N" else (k, m), dt)
                    B = O.uniform(rng, (k, n) if ob == "N" else (n, k), dt)
                    mb, nb, ka, kb = m // pr, n // pc, k // pc, k // pr
                    opA = A if oa == "N" else A.T            # index op(A) (m x k) / op(B) (k x n), then store pre-op
                    opB = B if ob == "N" else B.T
                    a_blk = opA[i * mb:(i + 1) * mb, j * ka:(j + 1) * ka]
                    b_blk = opB[i * kb:(i + 1) * kb, j * nb:(j + 1) * nb]
                    a_blk = np.ascontiguousarray(a_blk if oa == "N" else a_blk.T)
                    b_blk = np.ascontiguousarray(b_blk if ob == "N" else b_blk.T)
                    C = mg.gemm_2d(0.5, oa, torch.from_numpy(a_blk), ob, torch.from_numpy(b_blk), grid=(pr, pc),
                                   local_gemm=_local_gemm)
                    ref = O.gemm(0.5, oa, A, ob, B)[i * mb:(i + 1) * mb, j * nb:(j + 1) * nb]
                    assert tuple(C.shape) == ref.shape
                    res[f"gemm2d_{pr}x{pc}_{oa}{ob}_{dt}"] = float(np.linalg.norm(C.numpy() - ref)) / (
                        (k + 2) * 2.0 ** -53 * 0.5 * np.linalg.norm(A) * np.linalg.norm(B))
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_world4_gloo_2d_blocks():
    """mg.gemm_2d: 2 x 2, 1 x 4 and 4 x 1 grids, every op pair class, real and complex, blocks in stored layout."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker_2d, args=(r, 4, port, q)) for r in range(4)]
    for p in procs:
        p.start()
    out = dict(q.get(timeout=200) for _ in range(4))
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    for rank, res in out.items():
        assert len(res) == 24, (rank, len(res))
        for kname, err in res.items():
            assert err <= 2.0, (rank, kname, err)


def test_summa_steps_and_c_abi_partitioning():
    """The K-step schedule of mg.gemm_2d and the partitioning the C ABI exports (accblas_mg_row_panel,
    accblas_mg_k_chunks): pure host logic, no GPU."""
    import ctypes
    sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
    from accblas_b200 import _ffi
    from accblas_b200.mg import row_panels, summa_steps
    assert summa_steps(2, 2) == [(0, 0, 0, 0), (1, 0, 1, 0)]
    assert summa_steps(1, 4) == [(t, 0, 0, t) for t in range(4)]
    assert summa_steps(2, 3) == [(0, 0, 0, 0), (0, 1, 0, 1), (1, 0, 0, 2), (1, 1, 1, 0), (2, 0, 1, 1), (2, 1, 1, 2)]
    L = _ffi.lib()
    s, r = ctypes.c_int64(), ctypes.c_int64()
    for m, w in [(32768, 8), (300, 2), (5, 4), (1000, 3), (129, 8), (0, 2)]:
        mine = []
        for rk in range(w):
            assert L.accblas_mg_row_panel(m, w, rk, ctypes.byref(s), ctypes.byref(r)) == 0
            mine.append((s.value, r.value))
        assert mine == row_panels(m, w), (m, w, mine)          # the python planner and the library agree
    k0, dk = (ctypes.c_int64 * 64)(), (ctypes.c_int64 * 64)()
    for k in (0, 1, 100, 4095, 4096, 5000, 8192, 32768, 33000, 100000):
        n = L.accblas_mg_k_chunks(k, k0, dk, 64)
        ch = [(k0[i], dk[i]) for i in range(n)]
        assert sum(d for _, d in ch) == k and all(ch[i][0] + ch[i][1] == ch[i + 1][0] for i in range(n - 1)), (k, ch)
        assert all(c0 % 512 == 0 for c0, _ in ch) and (k < 4096 or ch[0][1] == 512), (k, ch)
    assert L.accblas_mg_k_chunks(1 << 20, k0, dk, 4) == 4 and k0[3] + dk[3] == 1 << 20   # cap: the last chunk takes the rest
    assert L.accblas_mg_row_panel(10, 0, 0, ctypes.byref(s), ctypes.byref(r)) != 0


def test_row_panels():
    sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))
    from accblas_b200.mg import row_panels
    assert row_panels(32768, 8) == [(i * 4096, 4096) for i in range(8)]
    p = row_panels(300, 2)
    assert p == [(0, 256), (256, 44)]
    p = row_panels(5, 4)
    assert p[0] == (0, 5) and all(r == 0 for _, r in p[1:])
    for m, w in [(1000, 3), (129, 8), (0, 2)]:
        p = row_panels(m, w)
        assert sum(r for _, r in p) == m and all(p[i][0] + p[i][1] == p[i + 1][0] or p[i + 1][1] == 0 for i in range(w - 1))


@pytest.mark.timeout(180)
def test_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = dict(q.get(timeout=150) for _ in range(2))
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    for rank, res in out.items():
        assert res, rank
        for kname, err in res.items():
            assert err <= 2.0, (rank, kname, err)
    assert any(k.startswith("gemm_") for k in out[0])
