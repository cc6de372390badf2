"""accblas_b200.mg -- one-process-per-GPU sharding of the hot path (torch.distributed plumbing).

The reference has no multi-GPU code; this layer is new and its parity oracle is
the single-GPU result.  Decomposition (BASELINE.json north_star / configs[4]):

  gemm   rows of op(A) and C are split into contiguous panels, one per rank; B is broadcast; C is
         gathered.  Operands start and the result ends on `root`.  The scatter of A, the local
         products and the gather of C are pipelined in sub-panels: while sub-panel i multiplies on
         the compute stream, sub-panel i+1 is arriving and sub-panel i-1 is leaving on the
         communicator's stream.
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


def row_panels(m: int, world: int, align: int = 128) -> List[Tuple[int, int]]:
    """Contiguous (start, rows) per rank; panel heights are multiples of `align` (the CTA tile) except
    possibly the last non-empty one.  Ranks past the end get (m, 0)."""
    per = -(-m // world)
    per = -(-per // align) * align
    out = []
    for r in range(world):
        s = min(r * per, m)
        out.append((s, min(per, m - s)))
    return out


def _default_gemm():
    import accblas_b200 as ab
    return ab.gemm


def gemm_rowsharded(alpha, opA, A: Optional[torch.Tensor], opB, B: Optional[torch.Tensor], *,
                    shapeA: Tuple[int, int], shapeB: Tuple[int, int], dtype: torch.dtype, device,
                    root: int = 0, group=None, subpanels: int = 4,
                    local_gemm: Optional[Callable] = None, out: Optional[torch.Tensor] = None):
    """C = alpha * op(A) * op(B) with rows of op(A)/C split over the ranks of `group`.

    A, B are given on `root` only (None elsewhere); shapeA/shapeB/dtype are known everywhere.
    Returns C on root (written into `out` if given), None on the other ranks.
    """
    gemm = local_gemm or _default_gemm()
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    opa = {"N": 0, "T": 1, "H": 2}.get(opA, opA)
    opb = {"N": 0, "T": 1, "H": 2}.get(opB, opB)
    m, k = (shapeA[0], shapeA[1]) if opa == 0 else (shapeA[1], shapeA[0])
    n = shapeB[1] if opb == 0 else shapeB[0]
    panels = row_panels(m, world)
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
        for i in range(depth):
            ops = []
            for r in remote:
                if i < len(rsubs[r]):
                    s, rows = rsubs[r][i]
                    ops.append(dist.P2POp(dist.irecv, C[s:s + rows], r, group))
            if ops:
                works += dist.batch_isend_irecv(ops)
        bwork.wait()
        for (s, rows) in subs(my_start, my_rows):
            gemm(alpha, opa, a_slice(s, rows), opb, B, out=C[s:s + rows])
        for w in works:
            w.wait()
        return C

    mine = subs(my_start, my_rows)
    bufs = [torch.empty(a_shape(rows), dtype=dtype, device=device) for (_, rows) in mine]
    rworks = [dist.irecv(b, src=root, group=group) for b in bufs]
    bwork.wait()
    sends, keep = [], []
    for (s, rows), buf, w in zip(mine, bufs, rworks):
        w.wait()
        c = gemm(alpha, opa, buf, opb, B)
        keep.append(c)
        sends.append(dist.isend(c, dst=root, group=group))
    for w in sends:
        w.wait()
    return None


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
        parts = [torch.empty_like(y) for _ in range(world)]
        dist.all_gather(parts, y, group=group)
        return torch.cat(parts)
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
