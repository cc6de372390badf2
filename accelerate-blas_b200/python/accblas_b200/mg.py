"""accblas_b200.mg -- one-process-per-GPU sharding of the hot path (torch.distributed plumbing).

The reference has no multi-GPU code; this layer is new and its parity oracle is
the single-GPU result.  Decomposition (BASELINE.json north_star / configs[4]):

  gemm   rows of op(A) and C are split into contiguous panels, one per rank; B is broadcast; C is
         gathered.  Operands start and the result ends on `root`.  On CUDA the root's A and C are
         mapped into every rank through CUDA IPC: A panels are pulled by the copy engines and the
         gemm kernels store their C tiles straight into the root's HBM over NVLink (transport
         "peer": the gather is the product's own epilogue).  Transport "p2p" (any backend) moves
         the panels with grouped send/recv, pipelined in sub-panels against the products.
  gemv   op N: rows of A and y are split (A pre-sharded), x broadcast, y all-gathered.
         op T/H: rows of A and x are split, every rank forms a full-length partial y, all-reduce.
  dot    contiguous slices (pre-sharded), all-reduce of one scalar.

Every function takes `local_*` callables defaulting to the CUDA library, so the sharding and
collective logic can be exercised on CPU with the gloo backend (tests/test_mg_cpu.py).
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist


def row_panels(m: int, world: int, align: int = 128, root: int = 0, root_bonus: int = 0) -> List[Tuple[int, int]]:
    """Contiguous (start, rows) per rank; panel heights are multiples of `align` (the CTA tile) except
    possibly the last non-empty one.  Ranks past the end get (m, 0).  `root_bonus` extra rows go to
    `root`, which can start multiplying while the others still wait for B."""
    bonus = max(0, min(int(root_bonus), m)) // align * align if world > 1 else 0
    per = -(-(m - bonus) // world)
    per = max(-(-per // align) * align, 0)
    out, s = [], 0
    for r in range(world):
        rows = min(per + (bonus if r == root else 0), m - s)
        out.append((s, rows))
        s += rows
    return out


def _default_gemm():
    import accblas_b200 as ab
    return ab.gemm


_ES = {torch.float32: 4, torch.float64: 8, torch.complex64: 8, torch.complex128: 16}
_CODE = {torch.float32: 0, torch.float64: 1, torch.complex64: 2, torch.complex128: 3}
# planning constants for the root bonus (measured on this pool: profiles/): real flop rate of the local
# gemm and the rate at which NCCL delivers B to the other ranks
_FLOPS = {torch.float32: 230e12, torch.float64: 36e12, torch.complex64: 228e12, torch.complex128: 36e12}
_BCAST_BPS = 670e9   # 8.6 GB in 12.8 ms at 8 ranks, 12.3 ms at 2 (profiles/r01_bench_n8_peer.json)


def root_bonus_rows(m: int, n: int, k: int, dtype, world: int) -> int:
    """Rows of head start for the root: what it multiplies while B is still on its way to the others."""
    if world <= 1 or dtype not in _FLOPS:
        return 0
    t_b = _ES[dtype] * k * n / _BCAST_BPS
    flops_row = (8.0 if dtype.is_complex else 2.0) * n * k
    rows = t_b * _FLOPS[dtype] / flops_row * (world - 1) / world
    return int(min(rows, m // (2 * world))) // 128 * 128


def gemm_rowsharded(alpha, opA, A: Optional[torch.Tensor], opB, B: Optional[torch.Tensor], *,
                    shapeA: Tuple[int, int], shapeB: Tuple[int, int], dtype: torch.dtype, device,
                    root: int = 0, group=None, subpanels: int = 4,
                    local_gemm: Optional[Callable] = None, out: Optional[torch.Tensor] = None,
                    transport: str = "auto", root_bonus: Optional[int] = None):
    """C = alpha * op(A) * op(B) with rows of op(A)/C split over the ranks of `group`.

    A, B are given on `root` only (None elsewhere); shapeA/shapeB/dtype are known everywhere.
    Returns C on root (written into `out` if given), None on the other ranks.

    transport = "peer" (default on CUDA): the other ranks map the root's A and C through CUDA IPC
        (accblas_peer_*): A row panels are pulled by the copy engines, and each rank's gemm kernel
        stores its C tiles directly into the root's HBM over NVLink -- the gather is fused into the
        product's epilogue.  NCCL carries only the broadcast of B and the completion handshake.
    transport = "p2p": scatter / gather with grouped send/recv on the communicator, pipelined in
        sub-panels (works on any backend; this is what the gloo CPU test drives).
    """
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    opa = {"N": 0, "T": 1, "H": 2}.get(opA, opA)
    opb = {"N": 0, "T": 1, "H": 2}.get(opB, opB)
    m, k = (shapeA[0], shapeA[1]) if opa == 0 else (shapeA[1], shapeA[0])
    n = shapeB[1] if opb == 0 else shapeB[0]
    if transport == "auto":
        transport = "peer" if (torch.device(device).type == "cuda" and local_gemm is None) else "p2p"
    if transport == "peer":
        if root_bonus is None:
            root_bonus = root_bonus_rows(m, n, k, dtype, world)
        return _gemm_rowsharded_peer(alpha, opa, A, opb, B, shapeA, shapeB, (m, n, k), dtype, device, root, group,
                                     rank, world, out, root_bonus)
    gemm = local_gemm or _default_gemm()
    panels = row_panels(m, world, root=root, root_bonus=root_bonus or 0)
    my_start, my_rows = panels[rank]

    # ---- B: broadcast (every rank needs all of it) ----
    if rank != root:
        B = torch.empty(shapeB, dtype=dtype, device=device)
    bwork = dist.broadcast(B, src=root, group=group, async_op=True)

    # ---- A: root keeps its panel and sends the others, sub-panel by sub-panel ----
    def subs(start, rows):
        if rows == 0:
            return []
        step = -(-rows // max(1, subpanels))
        step = -(-step // 128) * 128
        return [(start + o, min(step, rows - o)) for o in range(0, rows, step)]

    def a_slice(s, r):       # rows [s, s+r) of op(A) as a stored-layout contiguous tensor (root only)
        return A[s:s + r] if opa == 0 else A[:, s:s + r].contiguous()

    def a_shape(r):
        return (r, k) if opa == 0 else (k, r)

    if rank == root:
        C = out if out is not None else torch.empty((m, n), dtype=dtype, device=device)
        remote = [r for r in range(world) if r != root]
        rsubs = {r: subs(*panels[r]) for r in remote}
        depth = max([len(v) for v in rsubs.values()] + [0])
        keep, works = [], []
        # sub-panel-major order: sub-panel i of EVERY rank moves as one grouped batch, so all ranks
        # receive (and later return) their i-th piece at the same time instead of rank after rank
        for i in range(depth):
            ops = []
            for r in remote:
                if i < len(rsubs[r]):
                    t = a_slice(*rsubs[r][i])
                    keep.append(t)
                    ops.append(dist.P2POp(dist.isend, t, r, group))
            if ops:
                works += dist.batch_isend_irecv(ops)
        bwork.wait()
        mine = subs(my_start, my_rows)
        # The receive of sub-panel i is posted AFTER the root's own i-th product: the communicator's
        # stream waits for it, so the receive kernel starts when the others are about to deliver
        # instead of spinning on SMs the root's products need.
        for i in range(max(depth, len(mine))):
            if i < len(mine):
                s, rows = mine[i]
                gemm(alpha, opa, a_slice(s, rows), opb, B, out=C[s:s + rows])
            ops = []
            for r in remote:
                if i < len(rsubs[r]):
                    s, rows = rsubs[r][i]
                    ops.append(dist.P2POp(dist.irecv, C[s:s + rows], r, group))
            if ops:
                works += dist.batch_isend_irecv(ops)
        for w in works:
            w.wait()
        return C

    mine = subs(my_start, my_rows)
    bufs = [torch.empty(a_shape(rows), dtype=dtype, device=device) for (_, rows) in mine]
    rworks = [dist.batch_isend_irecv([dist.P2POp(dist.irecv, b, root, group)])[0] for b in bufs]
    bwork.wait()
    sends, keep = [], []
    for (s, rows), buf, w in zip(mine, bufs, rworks):
        w.wait()
        c = gemm(alpha, opa, buf, opb, B)
        keep.append(c)
        sends += dist.batch_isend_irecv([dist.P2POp(dist.isend, c, root, group)])
    for w in sends:
        w.wait()
    return None


def _gemm_rowsharded_peer(alpha, opa, A, opb, B, shapeA, shapeB, mnk, dtype, device, root, group, rank, world,
                          out, root_bonus):
    """transport = "peer": IPC-mapped root operands, copy-engine scatter, gather fused into the epilogue."""
    import ctypes
    import numpy as np
    from . import _ffi
    L = _ffi.lib()
    m, n, k = mnk
    es, code = _ES[dtype], _CODE[dtype]
    device = torch.device(device)
    panels = row_panels(m, world, root=root, root_bonus=root_bonus)
    start, rows = panels[rank]
    hsz = ctypes.sizeof(_ffi.PeerHandle)
    alpha_buf = np.array([alpha], dtype={0: np.float32, 1: np.float64, 2: np.complex64, 3: np.complex128}[code])
    al = alpha_buf.ctypes.data_as(ctypes.c_void_p)

    def stream():
        return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)

    def local_product(a_ptr, a_rows, c_ptr):
        rA, cA = (a_rows, k) if opa == 0 else (k, a_rows)
        st = L.accblas_gemm(stream(), code, opa, opb, rA, cA, shapeB[0], shapeB[1], al,
                            ctypes.c_void_p(a_ptr), ctypes.c_void_p(B.data_ptr()), ctypes.c_void_p(c_ptr))
        _ffi.check(st, "accblas_gemm (row panel)")

    with torch.cuda.device(device):
        # ---- handles of the root's A and C: 2 x 88 bytes, every call (the mapping itself is cached) ----
        if rank == root:
            if not (A.is_contiguous() and B.is_contiguous()):
                raise ValueError("gemm_rowsharded: operands must be dense row-major (as Accelerate's are)")
            C = out if out is not None else torch.empty((m, n), dtype=dtype, device=device)
            hs = (_ffi.PeerHandle * 2)()
            for i, t in enumerate((A, C)):
                st = L.accblas_peer_export(ctypes.c_void_p(t.data_ptr()), ctypes.byref(hs[i]))
                if st != 0:
                    raise _ffi.AccblasError(st, "accblas_peer_export",
                                            "operand is not a cudaMalloc allocation (expandable segments / "
                                            "cudaMallocAsync backend?): use transport='p2p'")
            hbuf = torch.frombuffer(bytearray(bytes(hs)), dtype=torch.uint8).to(device)
        else:
            hbuf = torch.empty(2 * hsz, dtype=torch.uint8, device=device)
            B = torch.empty(shapeB, dtype=dtype, device=device)
        dist.broadcast(hbuf, src=root, group=group)
        bwork = dist.broadcast(B, src=root, group=group, async_op=True)

        if rank == root:
            # the root multiplies its own panel while B is still travelling to the others
            if rows:
                a_off = start * k * es if opa == 0 else start * es
                if opa == 0:
                    local_product(A.data_ptr() + a_off, rows, C.data_ptr() + start * n * es)
                else:   # column slab of the stored matrix: densify locally (tight leading dimension)
                    slab = A[:, start:start + rows].contiguous()
                    local_product(slab.data_ptr(), rows, C.data_ptr() + start * n * es)
            bwork.wait()
        else:
            raw = bytes(hbuf.cpu().numpy().tobytes())
            hs = (_ffi.PeerHandle * 2).from_buffer_copy(raw)
            mapped = []
            for i in range(2):
                p = ctypes.c_void_p()
                _ffi.check(L.accblas_peer_open(ctypes.byref(hs[i]), ctypes.byref(p)), "accblas_peer_open")
                mapped.append(p.value)
            a_root, c_root = mapped
            if rows:
                # scatter: pull this rank's panel of op(A) with the copy engines (overlaps the broadcast)
                a_loc = torch.empty((rows, k) if opa == 0 else (k, rows), dtype=dtype, device=device)
                if opa == 0:
                    st = L.accblas_peer_copy2d(stream(), ctypes.c_void_p(a_loc.data_ptr()), rows * k * es,
                                               ctypes.c_void_p(a_root + start * k * es), rows * k * es,
                                               rows * k * es, 1)
                else:
                    st = L.accblas_peer_copy2d(stream(), ctypes.c_void_p(a_loc.data_ptr()), rows * es,
                                               ctypes.c_void_p(a_root + start * es), m * es, rows * es, k)
                _ffi.check(st, "accblas_peer_copy2d")
                bwork.wait()
                # product + gather in one kernel: C tiles are stored into the root's HBM
                local_product(a_loc.data_ptr(), rows, c_root + start * n * es)
            else:
                bwork.wait()
        # completion handshake, stream ordered on every rank: when it has run on the root, every
        # peer's kernel (and with it its remote stores) has finished
        done = torch.zeros(1, dtype=torch.float32, device=device)
        dist.all_reduce(done, group=group)
    return C if rank == root else None


def gemv_sharded(alpha, opA, A_local: torch.Tensor, x: torch.Tensor, *, group=None,
                 local_gemv: Optional[Callable] = None):
    """y = alpha * op(A) * x with the ROWS of the stored A pre-sharded (A_local = this rank's row panel).

    op N: x is the full vector on every rank (broadcast by the caller or replicated); returns the
          full y on every rank (all-gather of the row slices; panels must be equal-sized).
    op T/H: x is this rank's slice of the vector; returns the full y on every rank (all-reduce).
    """
    if local_gemv is None:
        import accblas_b200 as ab
        local_gemv = ab.gemv
    opa = {"N": 0, "T": 1, "H": 2}.get(opA, opA)
    world = dist.get_world_size(group)
    y = local_gemv(alpha, opa, A_local, x)
    if opa == 0:
        full = torch.empty((world * y.shape[0],), dtype=y.dtype, device=y.device)
        dist.all_gather_into_tensor(full, y, group=group)      # one collective, no list / cat copies
        return full
    if torch.is_complex(y):
        yr = torch.view_as_real(y)
        dist.all_reduce(yr, op=dist.ReduceOp.SUM, group=group)
        return y
    dist.all_reduce(y, op=dist.ReduceOp.SUM, group=group)
    return y


def dot_sharded(x_local: torch.Tensor, y_local: torch.Tensor, *, conj: bool = False, group=None,
                local_dot: Optional[Callable] = None):
    """dotu / dotc over pre-sharded slices: local reduction + all-reduce of one scalar."""
    if local_dot is None:
        import accblas_b200 as ab
        local_dot = ab.dotc if conj else ab.dotu
    d = local_dot(x_local, y_local).reshape(1)
    if torch.is_complex(d):
        dr = torch.view_as_real(d)
        dist.all_reduce(dr, op=dist.ReduceOp.SUM, group=group)
    else:
        dist.all_reduce(d, op=dist.ReduceOp.SUM, group=group)
    return d.reshape(())


# ----------------------------------------------------------------------------------------------------------
# 2-D block sharding: for products whose operands do not fit one GPU.
# The ranks form a pr x pc grid (rank = i * pc + j).  Rank (i, j) holds
#   block (i, j) of op(A): rows [i * m/pr, ...) x K-columns [j * k/pc, ...)     in its STORED (pre-op) layout,
#   block (i, j) of op(B): K-rows [i * k/pr, ...) x columns [j * n/pc, ...)     in its stored layout,
# and receives block (i, j) of C.  C_ij = op(A)_(i,:) * op(B)_(:,j): the A blocks are all-gathered along the
# process ROW, the B blocks along the process COLUMN, then ONE local product runs -- the library has no
# beta = 1 accumulation (the reference never needs it: beta = 0, LLVM/PTX/Level3.hs:90), so instead of SUMMA's
# K-step loop the whole K extent of the row / column panel is assembled first.  Per-rank memory is
# m k / pr + k n / pc + m n / (pr pc) elements; the bytes moved are SUMMA's.
# ----------------------------------------------------------------------------------------------------------
_GRID_GROUPS = {}


def _grid_groups(pr: int, pc: int, group=None):
    """Row and column process groups of the pr x pc grid (every rank creates every group, as torch requires)."""
    key = (pr, pc, id(group))
    if key not in _GRID_GROUPS:
        base = list(range(pr * pc)) if group is None else dist.get_process_group_ranks(group)
        rows = [dist.new_group([base[i * pc + j] for j in range(pc)]) for i in range(pr)]
        cols = [dist.new_group([base[i * pc + j] for i in range(pr)]) for j in range(pc)]
        _GRID_GROUPS[key] = (rows, cols)
    return _GRID_GROUPS[key]


def gemm_2d(alpha, opA, A_blk: torch.Tensor, opB, B_blk: torch.Tensor, *, grid: Tuple[int, int], group=None,
            local_gemm: Optional[Callable] = None) -> torch.Tensor:
    """Block (i, j) of C = alpha * op(A) * op(B) on a pr x pc grid of ranks (see above); blocks must be equal-sized."""
    pr, pc = grid
    world = dist.get_world_size(group)
    if pr * pc != world:
        raise ValueError(f"gemm_2d: grid {pr} x {pc} does not match the {world} ranks")
    opa = {"N": 0, "T": 1, "H": 2}.get(opA, opA)
    opb = {"N": 0, "T": 1, "H": 2}.get(opB, opB)
    local_gemm = local_gemm or _default_gemm()
    rank = dist.get_rank(group)
    i, j = divmod(rank, pc)
    rows, cols = _grid_groups(pr, pc, group)
    A_blk, B_blk = A_blk.contiguous(), B_blk.contiguous()
    # op(A)_(i,:): K runs along the stored columns for N, along the stored rows for T / H
    parts = [torch.empty_like(A_blk) for _ in range(pc)]
    dist.all_gather(parts, A_blk, group=rows[i])
    A_panel = torch.cat(parts, dim=1 if opa == 0 else 0) if pc > 1 else A_blk
    # op(B)_(:,j): K runs along the stored rows for N, along the stored columns for T / H
    parts = [torch.empty_like(B_blk) for _ in range(pr)]
    dist.all_gather(parts, B_blk, group=cols[j])
    B_panel = torch.cat(parts, dim=0 if opb == 0 else 1) if pr > 1 else B_blk
    return local_gemm(alpha, opa, A_panel, opb, B_panel)
