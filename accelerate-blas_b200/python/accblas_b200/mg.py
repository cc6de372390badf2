"""accblas_b200.mg -- one-process-per-GPU sharding of the hot path (torch.distributed plumbing).

The reference has no multi-GPU code; this layer is new and its parity oracle is
the single-GPU result.  Decomposition (BASELINE.json north_star / configs[4]):

  gemm   rows of op(A) and C are split into contiguous panels, one per rank; operands start and the
         result ends on `root`.  On CUDA (transport "peer") this module is a ctypes caller of the C
         ABI's accblas_mg_gemm / accblas_mg_gemm_plan (csrc/mg.cu): the root's A, B, C are mapped into
         every rank through CUDA IPC, A panels and B are pulled by the copy engines in K-chunks that
         overlap the chunk products, and the last chunk's kernel stores its C tiles straight into the
         root's HBM over NVLink.  Transport "p2p" (any backend) moves the panels with grouped
         send/recv, pipelined in sub-panels against the products (what the gloo CPU test drives).
  gemv   op N: rows of A and y are split (A pre-sharded), x broadcast, y all-gathered.
         op T/H: rows of A and x are split, every rank forms a full-length partial y, all-reduce.
  dot    contiguous slices (pre-sharded), all-reduce of one scalar.

Every function takes `local_*` callables defaulting to the CUDA library, so the sharding and
collective logic can be exercised on CPU with the gloo backend (tests/test_mg_cpu.py).
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import numpy as np
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
_NP = {torch.float32: np.float32, torch.float64: np.float64, torch.complex64: np.complex64, torch.complex128: np.complex128}
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


# ----------------------------------------------------------------------------------------------------------
# The C ABI's multi-GPU entry points (accblas_mg_*, csrc/mg.cu).  torch.distributed only ships the 128 bootstrap
# bytes; every transfer, token and exchange below happens inside libaccblas_b200.so.
# ----------------------------------------------------------------------------------------------------------
class Comm:
    """accblas_mg_comm_t of this process (one per process group, created collectively)."""

    def __init__(self, device, group=None):
        import ctypes
        from . import _ffi
        self._ffi, self._ct = _ffi, ctypes
        self.L = _ffi.lib()
        self.group = group
        self.device = torch.device(device)
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        ident = ctypes.create_string_buffer(128)
        if self.rank == 0:
            _ffi.check(self.L.accblas_mg_unique_id(ident), "accblas_mg_unique_id")
        box = [bytes(ident.raw)]
        src = 0 if group is None else dist.get_global_rank(group, 0)
        dist.broadcast_object_list(box, src=src, group=group)
        self.handle = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _ffi.check(self.L.accblas_mg_init(ctypes.byref(self.handle), self.rank, self.world, box[0]), "accblas_mg_init")

    def _stream(self):
        return self._ct.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        if self.handle:
            self.L.accblas_mg_finalize(self.handle)
            self.handle = None

    # ---- gemm ----
    def gemm_plan(self, opA, A, opB, B, *, shapeA, shapeB, dtype, out=None, root=0):
        return GemmPlan(self, opA, A, opB, B, shapeA, shapeB, dtype, out, root)

    # ---- gemv / dot over the exchange buffer ----
    def gemv(self, alpha, opA, A_local: torch.Tensor, x: torch.Tensor, out: Optional[torch.Tensor] = None,
             m_total: Optional[int] = None):
        """op N: the row panels may differ in height, down to empty (each rank's length travels with its deposit);
        ``m_total`` = rows of the whole matrix = length of y.  Without it the panels are taken to be equal."""
        ct = self._ct
        opa = {"N": 0, "T": 1, "H": 2}.get(opA, opA)
        A_local, x = A_local.contiguous(), x.contiguous()
        rows, cols = A_local.shape
        leny = (self.world * rows if m_total is None else int(m_total)) if opa == 0 else cols
        y = out if out is not None else torch.empty((leny,), dtype=A_local.dtype, device=A_local.device)
        al = np.array([alpha], dtype=_NP[A_local.dtype])
        with torch.cuda.device(self.device):
            st = self.L.accblas_mg_gemv(self.handle, self._stream(), _CODE[A_local.dtype], opa, rows, cols,
                                        al.ctypes.data_as(ct.c_void_p), ct.c_void_p(A_local.data_ptr()),
                                        ct.c_void_p(x.data_ptr()), x.shape[0], ct.c_void_p(y.data_ptr()))
        self._ffi.check(st, "accblas_mg_gemv")
        return y

    def dot(self, x: torch.Tensor, y: torch.Tensor, conj: bool = False):
        ct = self._ct
        x, y = x.contiguous(), y.contiguous()
        out = torch.empty((), dtype=x.dtype, device=x.device)
        with torch.cuda.device(self.device):
            st = self.L.accblas_mg_dot(self.handle, self._stream(), _CODE[x.dtype], int(conj), min(x.shape[0], y.shape[0]),
                                       ct.c_void_p(x.data_ptr()), ct.c_void_p(y.data_ptr()), ct.c_void_p(out.data_ptr()))
        self._ffi.check(st, "accblas_mg_dot")
        return out


class GemmPlan:
    """accblas_mg_plan_t: the root's A, B, C mapped into every rank, local replicas allocated; run() is the product."""

    def __init__(self, comm: Comm, opA, A, opB, B, shapeA, shapeB, dtype, out, root):
        ct = comm._ct
        self.comm = comm
        self.opa = {"N": 0, "T": 1, "H": 2}.get(opA, opA)
        self.opb = {"N": 0, "T": 1, "H": 2}.get(opB, opB)
        m = shapeA[0] if self.opa == 0 else shapeA[1]
        n = shapeB[1] if self.opb == 0 else shapeB[0]
        self.dtype, self.root = dtype, root
        self.C = None
        ptr = lambda t: ct.c_void_p(t.data_ptr() if t is not None else 0)  # noqa: E731
        if comm.rank == root:
            if not (A.is_contiguous() and B.is_contiguous()):
                raise ValueError("gemm plan: operands must be dense row-major (as Accelerate's are)")
            self.C = out if out is not None else torch.empty((m, n), dtype=dtype, device=comm.device)
            if not self.C.is_contiguous():
                raise ValueError("gemm plan: out must be contiguous")
            self._keep = (A, B, self.C)
        else:
            A = B = None
        self.handle = ct.c_void_p()
        with torch.cuda.device(comm.device):
            st = comm.L.accblas_mg_gemm_plan(comm.handle, _CODE[dtype], self.opa, self.opb, shapeA[0], shapeA[1], shapeB[0],
                                             shapeB[1], ptr(A), ptr(B), ptr(self.C), root, ct.byref(self.handle))
        comm._ffi.check(st, "accblas_mg_gemm_plan")

    def run(self, alpha=1.0):
        ct = self.comm._ct
        al = np.array([alpha], dtype=_NP[self.dtype])
        with torch.cuda.device(self.comm.device):
            st = self.comm.L.accblas_mg_gemm_run(self.handle, self.comm._stream(), al.ctypes.data_as(ct.c_void_p))
        self.comm._ffi.check(st, "accblas_mg_gemm_run")
        return self.C

    def destroy(self):
        if self.handle:
            self.comm.L.accblas_mg_gemm_plan_destroy(self.handle)
            self.handle = None


_COMMS = {}


def get_comm(device, group=None) -> Comm:
    """The process's communicator for `group` (created collectively on first use, then cached)."""
    key = (id(group), torch.device(device).index)
    if key not in _COMMS:
        _COMMS[key] = Comm(device, group)
    return _COMMS[key]


def _gemm_rowsharded_peer(alpha, opa, A, opb, B, shapeA, shapeB, mnk, dtype, device, root, group, rank, world,
                          out, root_bonus):
    """transport = "peer": one accblas_mg_gemm (plan + run + destroy) through the C ABI."""
    comm = get_comm(device, group)
    plan = comm.gemm_plan(opa, A, opb, B, shapeA=shapeA, shapeB=shapeB, dtype=dtype, out=out, root=root)
    try:
        C = plan.run(alpha)
        torch.cuda.current_stream(torch.device(device)).synchronize()
    finally:
        plan.destroy()
    return C if rank == root else None


def gemv_sharded(alpha, opA, A_local: torch.Tensor, x: torch.Tensor, *, group=None,
                 local_gemv: Optional[Callable] = None, m_total: Optional[int] = None):
    """y = alpha * op(A) * x with the ROWS of the stored A pre-sharded (A_local = this rank's row panel).

    op N: x is the full vector on every rank (broadcast by the caller or replicated); returns the
          full y on every rank (all-gather of the row slices).  Panels cut by ``row_panels(m_total, world)``
          (= accblas_mg_row_panel: whole 128-row tiles, so the last ones can be short or empty) need ``m_total``;
          without it the panels must be equal-sized.
    op T/H: x is this rank's slice of the vector; returns the full y on every rank (all-reduce).
    """
    if local_gemv is None and A_local.is_cuda:
        return get_comm(A_local.device, group).gemv(alpha, opA, A_local, x, m_total=m_total)   # exchange over peer memory (C ABI)
    if local_gemv is None:
        import accblas_b200 as ab
        local_gemv = ab.gemv
    opa = {"N": 0, "T": 1, "H": 2}.get(opA, opA)
    world = dist.get_world_size(group)
    if opa == 0 and m_total is not None:
        # ragged panels on a backend without a variable-length gather: pad every slice to the tallest panel
        panels = row_panels(int(m_total), world)
        rows = panels[dist.get_rank(group)][1]
        if rows != A_local.shape[0]:
            raise ValueError(f"gemv_sharded: this rank's panel has {A_local.shape[0]} rows, row_panels gives {rows}")
        tall = max(r for _, r in panels)
        padded = torch.zeros((tall,), dtype=A_local.dtype, device=A_local.device)
        if rows > 0:
            padded[:rows] = local_gemv(alpha, opa, A_local, x)
        full = torch.empty((world * tall,), dtype=padded.dtype, device=padded.device)
        dist.all_gather_into_tensor(full, padded, group=group)
        return torch.cat([full[r * tall:r * tall + n] for r, (_, n) in enumerate(panels)])
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
    if local_dot is None and x_local.is_cuda:
        return get_comm(x_local.device, group).dot(x_local, y_local, conj)        # exchange over peer memory (C ABI)
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
# 2-D block sharding: SUMMA, for products whose operands do not fit one GPU.
# The ranks form a pr x pc grid (rank = i * pc + j).  Rank (i, j) holds
#   block (i, j) of op(A): rows [i * m/pr, ...) x K-columns [j * k/pc, ...)     in its STORED (pre-op) layout,
#   block (i, j) of op(B): K-rows [i * k/pr, ...) x columns [j * n/pc, ...)     in its stored layout,
# and receives block (i, j) of C.  K is cut into L = lcm(pr, pc) steps of k / L (the common refinement of the two
# block distributions).  Step t: the process-ROW broadcasts the K-slab t of op(A)_(i,:) from the rank that owns it,
# the process-COLUMN broadcasts the K-slab t of op(B)_(:,j), and every rank accumulates
#       C_ij = alpha * slab_A * slab_B + C_ij          (beta = 1 in the product's epilogue: accblas_gemm_ex)
# The broadcasts of step t + 1 are in flight while step t multiplies (two slab buffers per operand).
# Per-rank memory: the three blocks + two K-slabs of each operand -- never a whole row / column panel.
# ----------------------------------------------------------------------------------------------------------
_GRID_GROUPS = {}


def _grid_groups(pr: int, pc: int, group=None):
    """Row and column process groups of the pr x pc grid (every rank creates every group, as torch requires)."""
    key = (pr, pc, id(group))
    if key not in _GRID_GROUPS:
        base = list(range(pr * pc)) if group is None else dist.get_process_group_ranks(group)
        rows = [dist.new_group([base[i * pc + j] for j in range(pc)]) for i in range(pr)]
        cols = [dist.new_group([base[i * pc + j] for i in range(pr)]) for j in range(pc)]
        _GRID_GROUPS[key] = (rows, cols, base)
    return _GRID_GROUPS[key]


def summa_steps(pr: int, pc: int):
    """[(owner column of the A slab, slab index inside that block, owner row of the B slab, slab index inside it)]
    for the L = lcm(pr, pc) K-steps."""
    from math import gcd
    L = pr * pc // gcd(pr, pc)
    return [(t // (L // pc), t % (L // pc), t // (L // pr), t % (L // pr)) for t in range(L)]


def gemm_2d(alpha, opA, A_blk: torch.Tensor, opB, B_blk: torch.Tensor, *, grid: Tuple[int, int], group=None,
            local_gemm: Optional[Callable] = None) -> torch.Tensor:
    """Block (i, j) of C = alpha * op(A) * op(B) on a pr x pc grid of ranks (SUMMA, see above); blocks must be
    equal-sized and the K extent of a block divisible by lcm(pr, pc) / pc (resp. / pr).
    local_gemm(alpha, opa, A, opb, B, out=, beta=, Cin=) defaults to accblas_b200.gemm."""
    pr, pc = grid
    world = dist.get_world_size(group)
    if pr * pc != world:
        raise ValueError(f"gemm_2d: grid {pr} x {pc} does not match the {world} ranks")
    opa = {"N": 0, "T": 1, "H": 2}.get(opA, opA)
    opb = {"N": 0, "T": 1, "H": 2}.get(opB, opB)
    local_gemm = local_gemm or _default_gemm()
    rank = dist.get_rank(group)
    i, j = divmod(rank, pc)
    rows, cols, base = _grid_groups(pr, pc, group)
    A_blk, B_blk = A_blk.contiguous(), B_blk.contiguous()
    steps = summa_steps(pr, pc)
    L = len(steps)
    # K extents: op(A) block is mb x ka with ka = k / pc; op(B) block is kb x nb with kb = k / pr; a step moves kk = k / L
    ka = A_blk.shape[1] if opa == 0 else A_blk.shape[0]
    kb = B_blk.shape[0] if opb == 0 else B_blk.shape[1]
    mb = A_blk.shape[0] if opa == 0 else A_blk.shape[1]
    nb = B_blk.shape[1] if opb == 0 else B_blk.shape[0]
    if ka * pc != kb * pr or (ka * pc) % L:
        raise ValueError("gemm_2d: the K extents of the blocks do not tile the common refinement")
    kk = ka * pc // L

    def a_slab(s):      # K-slab s of my op(A) block, in stored layout, contiguous
        return (A_blk[:, s * kk:(s + 1) * kk] if opa == 0 else A_blk[s * kk:(s + 1) * kk]).contiguous()

    def b_slab(s):
        return (B_blk[s * kk:(s + 1) * kk] if opb == 0 else B_blk[:, s * kk:(s + 1) * kk]).contiguous()

    a_shape = (mb, kk) if opa == 0 else (kk, mb)
    b_shape = (kk, nb) if opb == 0 else (nb, kk)

    def post(t):        # start the two broadcasts of step t; returns (A slab, B slab, works)
        ja, sa, ib, sb = steps[t]
        abuf = a_slab(sa) if ja == j else torch.empty(a_shape, dtype=A_blk.dtype, device=A_blk.device)
        bbuf = b_slab(sb) if ib == i else torch.empty(b_shape, dtype=B_blk.dtype, device=B_blk.device)
        works = []
        if pc > 1:
            works.append(dist.broadcast(abuf, src=base[i * pc + ja], group=rows[i], async_op=True))
        if pr > 1:
            works.append(dist.broadcast(bbuf, src=base[ib * pc + j], group=cols[j], async_op=True))
        return abuf, bbuf, works

    C = torch.empty((mb, nb), dtype=A_blk.dtype, device=A_blk.device)
    nxt = post(0)
    for t in range(L):
        abuf, bbuf, works = nxt
        if t + 1 < L:
            nxt = post(t + 1)                       # step t + 1 travels while step t multiplies
        for w in works:
            w.wait()
        if t == 0:
            local_gemm(alpha, opa, abuf, opb, bbuf, out=C)
        else:
            local_gemm(alpha, opa, abuf, opb, bbuf, out=C, beta=1, Cin=C)
    return C
