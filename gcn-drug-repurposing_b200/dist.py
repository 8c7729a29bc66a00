"""Row-partitioned SpMM for graphs that exceed one GPU (north_star: R-MAT 16M nodes / 512M edges over the
8 B200s of one box).  One process per GPU (torch.distributed, NCCL over NVLink 5 / NVSwitch).

Partition: p contiguous row blocks of Â balanced by nnz (R-MAT is skewed); rank r owns rows [lo_r, hi_r) of
Â, of X and of Y; output rows stay local (they are the next layer's X block).  Âᵀ·G (backward) uses a second
partition of Âᵀ built the same way, so accumulation order stays deterministic (no reduce-scatter).

Exchange step.  A rank does not need all of X: it needs the rows named by the column ids of its block.  On
R-MAT graphs that is ~25-35 % of the remote rows at p = 8 (measured), so instead of an
all-gather (7/8 of X inbound per rank) every rank
    1. packs, per destination, the rows of its X block that destination references (gdr_gather_rows_f32),
    2. exchanges them with ONE NCCL all-to-all (variable split sizes, lists fixed at partition time),
    3. multiplies: the column ids of the local block are renumbered once into the compact buffer
       [own rows | rows received from rank 0 | rank 1 | ...].
Overlap: the local block is split by column owner into Â_own (columns in the own block) and Â_rem (the
rest).  Â_own·X_own runs on the compute stream while pack + all-to-all run on a side stream; Â_rem·X_recv
then accumulates in place (gdr_spmm_csr_f32 with e1 = y).  Results are deterministic; they differ from the
1-GPU product only in fp32 summation order (own-block terms first).

The reference has no distributed code at all; this module is new surface, kept to the one
exchange step the path really has.

Classes, in the order they were measured (DESIGN.md §4 has the numbers):
    RowPartitionedSpMM   the design above (pull only, NCCL all-to-all)
    LShapedSpMM          blocks above the diagonal are multiplied at the column owner and their partial rows
                         sent back; transport "nccl" or "p2p" (gdr_push_rows_f32 into peer symmetric memory);
                         options hub_first / split_pull reorder the step around rank 0's link (the critical path)
(An NVSwitch-multicast variant for the hub ranks - multimem.st broadcast of their X block, multimem.ld_reduce of
their partial rows - was measured in round 1 at 8.55 ms against 7.28 ms per step and removed; the two multimem
kernels stay in csrc/p2p.cu as tested primitives.)
"""
import numpy as np
import torch
import torch.distributed as dist

from . import _ffi
from .graph import DeviceCSR


def nnz_balanced_bounds(rowptr, parts):
    """Row boundaries [parts+1] of contiguous blocks with ~equal nnz.  rowptr: numpy int array [n+1]."""
    rowptr = np.asarray(rowptr)
    nnz = int(rowptr[-1])
    targets = (np.arange(1, parts) * (nnz / parts)).astype(np.int64)
    cuts = np.searchsorted(rowptr, targets, side="left")
    b = np.concatenate([[0], cuts, [len(rowptr) - 1]]).astype(np.int64)
    return np.maximum.accumulate(b)


def plan_exchange(col, bounds, rank):
    """Host/device-agnostic index arithmetic of the compressed exchange (pure torch ops, unit-tested on CPU).
    col: int64 global column ids of this rank's block; bounds: int64 [p+1].
    Returns (need_from[p] lists of owner-local row ids this rank needs from each owner (own entry empty),
             new_col int64: column ids renumbered into [own block | received rows in owner order],
             is_own bool mask per entry)."""
    p = len(bounds) - 1
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    is_own = (col >= lo) & (col < hi)
    remote = torch.unique(col[~is_own])                              # sorted => grouped by owner
    b = torch.as_tensor(bounds, dtype=torch.int64, device=col.device)
    cut = torch.searchsorted(remote, b)                              # [p+1] positions of the owner groups
    need_from = [remote[int(cut[s]):int(cut[s + 1])] - int(bounds[s]) for s in range(p)]
    pos = torch.searchsorted(remote, col)
    new_col = torch.where(is_own, col - lo, (hi - lo) + pos)
    return need_from, new_col, is_own


def split_csr(rowptr, new_col, val, keep, n_rows, n_cols):
    """Sub-matrix of the entries with keep == True (row order and in-row order preserved)."""
    deg = (rowptr[1:] - rowptr[:-1]).to(torch.int64)
    row_of = torch.repeat_interleave(torch.arange(n_rows, device=rowptr.device), deg)
    cnt = torch.bincount(row_of[keep], minlength=n_rows)
    rp = torch.zeros(n_rows + 1, dtype=torch.int64, device=rowptr.device)
    rp[1:] = torch.cumsum(cnt, 0)
    return DeviceCSR(rp.to(torch.int32), new_col[keep].to(torch.int32).contiguous(), val[keep].contiguous(),
                     n_rows, n_cols)


def p2p_available(device, group=None):
    """True on every rank iff a symmetric-memory rendezvous (peer pointers over NVLink) comes up on all of them.
    Collective.  Used to choose between the two CUDA transports of LShapedSpMM ("p2p" stores vs NCCL all-to-all);
    it never selects a CPU path."""
    ok = 1
    try:
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(1024, dtype=torch.float32, device=device)
        h = symm.rendezvous(t, (group or dist.group.WORLD).group_name)
        ok = 1 if len(h.buffer_ptrs) == dist.get_world_size(group) and all(int(p) != 0 for p in h.buffer_ptrs) else 0
    except Exception:  # noqa: BLE001  (any failure means "use the NCCL transport")
        ok = 0
    flag = torch.tensor([ok], dtype=torch.int32, device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    return bool(flag.item())


class RowPartitionedSpMM:
    """This rank's row block of Â plus the buffers and index lists of the exchange step."""

    def __init__(self, local_csr, bounds, rank, world, d, group=None, overlap=True):
        self.bounds = [int(b) for b in bounds]
        self.rank, self.world, self.d, self.group = rank, world, d, group
        self.n = self.bounds[-1]
        dev = local_csr.device
        lo, hi = self.bounds[rank], self.bounds[rank + 1]
        self.n_own = hi - lo
        self.overlap = overlap and world > 1
        if world == 1:
            self.csr_all = local_csr
            self.xbuf = torch.empty(self.n_own, d, dtype=torch.float32, device=dev)
            self.x_own = self.xbuf
            self.y = torch.empty(self.n_own, d, dtype=torch.float32, device=dev)
            self.transport = "single GPU"
            self.recv_rows = 0
            return
        need_from, new_col, is_own = plan_exchange(local_csr.col.to(torch.int64), self.bounds, rank)
        recv_counts = torch.tensor([int(t.numel()) for t in need_from], dtype=torch.int64, device=dev)
        send_counts = torch.empty_like(recv_counts)
        dist.all_to_all_single(send_counts, recv_counts, group=group)          # who needs how many of my rows
        self.recv_splits = recv_counts.tolist()
        self.send_splits = send_counts.tolist()
        want = torch.cat(need_from) if sum(self.recv_splits) else torch.zeros(0, dtype=torch.int64, device=dev)
        self.send_idx = torch.empty(sum(self.send_splits), dtype=torch.int64, device=dev)
        dist.all_to_all_single(self.send_idx, want, self.send_splits, self.recv_splits, group=group)
        self.recv_rows = int(sum(self.recv_splits))
        n_buf = self.n_own + self.recv_rows
        self.xbuf = torch.empty(n_buf, d, dtype=torch.float32, device=dev)
        self.x_own = self.xbuf[:self.n_own]                  # callers write their X block here (no copy)
        self.x_recv = self.xbuf[self.n_own:]
        self.sendbuf = torch.empty(int(sum(self.send_splits)), d, dtype=torch.float32, device=dev)
        self.y = torch.empty(self.n_own, d, dtype=torch.float32, device=dev)
        if self.overlap:
            self.csr_own = split_csr(local_csr.rowptr, new_col, local_csr.val, is_own, self.n_own, n_buf)
            self.csr_rem = split_csr(local_csr.rowptr, new_col, local_csr.val, ~is_own, self.n_own, n_buf)
        else:
            self.csr_all = DeviceCSR(local_csr.rowptr, new_col.to(torch.int32).contiguous(), local_csr.val,
                                     self.n_own, n_buf)
        self.comm_stream = torch.cuda.Stream(device=dev) if dev.type == 'cuda' else None
        self.transport = ("nccl all_to_all of referenced rows only (%s)" %
                          ("overlapped with the own-column SpMM" if self.overlap else "then one local SpMM"))

    @classmethod
    def from_full(cls, full_csr, bounds, rank, world, d, group=None, overlap=True):
        lo, hi = int(bounds[rank]), int(bounds[rank + 1])
        p0, p1 = int(full_csr.rowptr[lo].item()), int(full_csr.rowptr[hi].item())
        local = DeviceCSR((full_csr.rowptr[lo:hi + 1] - p0).contiguous(), full_csr.col[p0:p1].clone(),
                          full_csr.val[p0:p1].clone(), hi - lo, full_csr.n_cols)
        return cls(local, bounds, rank, world, d, group, overlap)

    def _exchange(self):
        L = _ffi.lib()
        _ffi.check(L.gdr_gather_rows_f32(_ffi.ptr(self.x_own), _ffi.ptr(self.send_idx), self.send_idx.numel(),
                                         self.n_own, self.d, _ffi.ptr(self.sendbuf), _ffi.stream()), "gather_rows")
        dist.all_to_all_single(self.x_recv, self.sendbuf, self.recv_splits, self.send_splits, group=self.group)

    def spmm(self, x_local=None):
        """Y_block = (Â·X)[lo:hi].  x_local: this rank's rows of X, or None when the caller has already
        written them into `self.x_own`."""
        if x_local is not None and x_local.data_ptr() != self.x_own.data_ptr():
            self.x_own.copy_(x_local)
        if self.world == 1:
            return self.csr_all.spmm(self.x_own, out=self.y)
        if not self.overlap:
            self._exchange()
            return self.csr_all.spmm(self.xbuf, out=self.y)
        cur = torch.cuda.current_stream()
        self.comm_stream.wait_stream(cur)
        with torch.cuda.stream(self.comm_stream):
            self._exchange()
        self.csr_own.spmm(self.xbuf, out=self.y)                 # touches only the own rows of xbuf
        cur.wait_stream(self.comm_stream)
        return self.csr_rem.spmm(self.xbuf, e1=self.y, out=self.y)


# ======================================================================================================
# L-shaped partition: rows stay with their owner, but a block whose COLUMN owner has the higher rank is
# multiplied there and its partial output rows are sent back ("push"), instead of pulling the X rows.
# ======================================================================================================
"""Why.  On skewed (R-MAT) graphs the row blocks of the low ranks are hub rows: rank 0's rows reference most
columns of every other rank (pull volume ~ |unique columns| per block), while the block has few rows, so
shipping the block's PARTIAL OUTPUT ROWS (|non-empty rows| per block) is several times cheaper.  With
contiguous nnz-ordered blocks the cheaper direction is, to a good approximation, "block (r, s) is computed
by rank max(r, s)":
    r > s  (lower triangle)  pull:  rank s sends the referenced rows of X_s to r,      r computes Â[r,s]·X_s
    r < s  (upper triangle)  push:  rank s computes Â[r,s]·X_s from its OWN X block,   sends the non-empty
                                    output rows to r, which adds them in fixed rank order (deterministic).
Rank s therefore owns an L-shaped region of Â (its row block left of the diagonal, plus its column block
above the diagonal); bounds are chosen so that the L-shapes, not the row blocks, carry equal cost.  Measured
on R-MAT scale 20, p = 8: bottleneck volume (max over ranks of rows sent or received) 496 K -> 323 K rows.

The second gain is the dependency structure: the push product needs only the rank's own X, so it runs while
the X rows travel; the partial rows travel while the pull product runs; only the final add waits.  No product
is split into extra passes over Y."""


def coarse_histogram(row_global, col_global, n, bins):
    """[bins, bins] int64 counts of entries per (row bin, column bin); bin = index // ceil(n / bins)."""
    size = -(-n // bins)
    key = (row_global // size) * bins + (col_global // size)
    return torch.bincount(key, minlength=bins * bins).reshape(bins, bins), size


def lshape_bounds(hist, bin_size, n, parts, row_weight=6.0, comm_weight=0.0, rank_weights=None):
    """Bounds [parts+1] (multiples of bin_size, last = n) that minimise the largest L-shape cost
        cost(s) = nnz(L_s) + row_weight * (own rows + expected non-empty rows of the push blocks computed here
                                           + add pass: own rows + expected partial rows received)
                           + comm_weight * max(rows sent, rows received)      (X rows and partial rows together)
    by bisection on the cost cap with a greedy left-to-right fill.  All row counts are Poisson estimates from the
    coarse histogram (rows of a bin hit by lambda entries each: non-empty fraction 1 - exp(-lambda)); for traffic
    towards the not-yet-placed higher ranks the remaining index range is taken as (parts - s - 1) equal chunks.
    comm_weight is in nnz-equivalents per row moved (a 512-byte row at ~750 GB/s costs ~14 gathers of ~50 ps).
    rank_weights [parts]: multiplier on each rank's modelled cost (default 1).  The model knows nothing about who waits
    for whom inside a step, so `rebalance_weights` feeds the MEASURED finish time of every rank back through it.
    hist: numpy [B, B] from coarse_histogram (summed over ranks)."""
    hist = np.asarray(hist, dtype=np.float64)
    nb = hist.shape[0]
    rw = np.ones(parts) if rank_weights is None else np.asarray(rank_weights, dtype=np.float64)
    sq = np.zeros(nb + 1)
    sq[1:] = np.diagonal(hist.cumsum(0).cumsum(1))                      # nnz of the leading square [0,b)x[0,b)
    ccum = np.zeros((nb, nb + 1))
    ccum[:, 1:] = hist.cumsum(1)                                         # per row bin, cumulative over columns
    rcum = np.zeros((nb + 1, nb))
    rcum[1:, :] = hist.cumsum(0)                                         # per column bin, cumulative over rows
    rows_in = np.minimum(np.arange(1, nb + 1) * bin_size, n) - np.minimum(np.arange(nb) * bin_size, n)
    rows_in = np.maximum(rows_in, 1).astype(np.float64)

    def hit(lam_total, rows, chunks=1):
        """expected (row, chunk) pairs with at least one entry, lam_total entries spread over rows x chunks"""
        return rows * chunks * (1.0 - np.exp(-lam_total / (rows * chunks)))

    def costs_from(b0, s):
        """cost of rank s's L-shape starting at bin b0, for every end bin b1 in (b0, nb]  -> array [nb - b0]."""
        w = nb - b0
        ends = np.arange(b0 + 1, nb + 1)
        c = sq[ends] - sq[b0]
        own = np.minimum(ends * bin_size, n) - min(b0 * bin_size, n)
        inside = np.triu(np.ones((w, w)))                                # [bin offset i, end offset e]: i < end
        if b0 > 0:
            lam = ccum[:b0, b0 + 1:nb + 1] - ccum[:b0, b0:b0 + 1]       # rows of lower ranks x columns [b0, end)
            push_rows = hit(lam, rows_in[:b0, None]).sum(0)              # partial rows computed here and sent down
            low = rcum[b0 + 1:nb + 1, :b0].T - rcum[b0:b0 + 1, :b0].T    # [col bin < b0, end]: rows [b0, end)
            recv_x = hit(low, rows_in[:b0, None]).sum(0)                 # X rows pulled from lower ranks
        else:
            push_rows, recv_x = 0.0, 0.0
        m = max(parts - s - 1, 1)
        right = ccum[b0:, nb:nb + 1] - ccum[b0:, b0 + 1:nb + 1]          # [row bin >= b0, end]: nnz in cols >= end
        recv_p = (hit(right, rows_in[b0:, None], m) * inside).sum(0)     # partial rows received from higher ranks
        recv_rows = (hit(right, rows_in[b0:, None]) * inside).sum(0)     # distinct own rows that receive any
        below = rcum[nb:nb + 1, b0:].T - rcum[b0 + 1:nb + 1, b0:].T      # [col bin >= b0, end]: nnz in rows >= end
        sent_x = (hit(below, rows_in[b0:, None], m) * inside).sum(0)     # X rows pushed up to higher ranks
        if s == parts - 1:
            recv_p, sent_x, recv_rows = 0.0 * recv_p, 0.0 * sent_x, 0.0 * recv_rows
        add = np.where(recv_rows > 0, own + recv_p, 0.0)
        comm = np.maximum(sent_x + push_rows, recv_x + recv_p)
        return np.maximum.accumulate(c + row_weight * (own + push_rows + add) + comm_weight * comm)

    def fill(cap):
        b, out = 0, [0]
        for s in range(parts):
            if b >= nb:
                out.append(nb)
                continue
            c = costs_from(b, s) * rw[s]
            k = int(np.searchsorted(c, cap, side="right"))               # number of end bins with cost <= cap
            b = b + max(k, 1)
            out.append(min(b, nb))
        return out

    lo, hi = 0.0, float(costs_from(0, 0)[-1]) * float(rw.max()) * 1.01 + 1.0
    for _ in range(30):
        mid = 0.5 * (lo + hi)
        if fill(mid)[-1] >= nb:
            hi = mid
        else:
            lo = mid
    b = fill(hi)
    b[-1] = nb
    return np.minimum(np.asarray(b, dtype=np.int64) * bin_size, n)


def rebalance_weights(weights, finish_ms, damping=0.8, clip=(0.25, 4.0)):
    """New rank_weights for lshape_bounds from the measured per-rank finish times of a step (time from the common
    start until the rank's Y block is complete, waits on other ranks' rows included): a rank that finishes late gets a
    proportionally heavier cost, hence a smaller region.  Normalised to mean 1."""
    w = np.asarray(weights, dtype=np.float64)
    t = np.asarray(finish_ms, dtype=np.float64)
    w = w * (t / t.mean()) ** damping
    w = np.clip(w / w.mean(), clip[0], clip[1])
    return w / w.mean()


def csr_from_sorted_rows(row, col, val, n_rows, n_cols):
    """CSR from entries already ordered by (row, then the wanted in-row order).  Pure torch (CPU or CUDA)."""
    rp = torch.zeros(n_rows + 1, dtype=torch.int64, device=row.device)
    if row.numel():
        rp[1:] = torch.cumsum(torch.bincount(row, minlength=n_rows), 0)
    return DeviceCSR(rp.to(torch.int32), col.to(torch.int32).contiguous(), val.contiguous(), n_rows, n_cols)


def _a2a_var(t, send_splits, group):
    """Variable all-to-all of a 1-D tensor: exchanges the counts first; returns (received, recv_splits)."""
    dev = t.device
    s = torch.tensor(send_splits, dtype=torch.int64, device=dev)
    r = torch.empty_like(s)
    dist.all_to_all_single(r, s, group=group)
    recv_splits = r.tolist()
    out = torch.empty(int(sum(recv_splits)), dtype=t.dtype, device=dev)
    dist.all_to_all_single(out, t.contiguous(), recv_splits, list(send_splits), group=group)
    return out, recv_splits


def cyclic_vblock_layout(vb, world):
    """Cyclic ownership of V = world * m VIRTUAL blocks.  vb: bounds [V+1] of the virtual blocks in the natural vertex
    order (lshape_bounds(..., parts=V): equal L-shape costs); virtual block k belongs to rank k % world.  Every rank then
    owns a hub-like block, m - 2 middle ones and a tail-like one, so the X rows and partial rows that the plain L-shape
    funnels through rank 0's NVLink port (2.2 GB each way per step on R-MAT-24) are spread over all ports.
    Vertices are renumbered so that a rank's blocks are contiguous (rank r: blocks r, r + world, r + 2 world, ...).
    Returns (order: virtual block id of the q-th block in the new numbering, new_vb: its bounds [V+1] in the new
    numbering, rank_bounds [world+1])."""
    vb = np.asarray(vb, dtype=np.int64)
    v = len(vb) - 1
    if v % world:
        raise ValueError(f"cyclic_vblock_layout: {v} virtual blocks do not divide over {world} ranks")
    m = v // world
    order = np.array([t * world + r for r in range(world) for t in range(m)], dtype=np.int64)
    sizes = vb[order + 1] - vb[order]
    new_vb = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    return order, new_vb, new_vb[::m].copy()


def cyclic_permutation(vb, order, device):
    """perm[new id] = old id for the layout above, and its inverse (int64 device tensors)."""
    perm = torch.cat([torch.arange(int(vb[k]), int(vb[k + 1]), device=device, dtype=torch.int64) for k in order])
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(perm.numel(), device=device, dtype=torch.int64)
    return perm, inv


class LShapedSpMM:
    """Rank s's share of Y = Â·X under the L-shaped partition (see the comment above).  Construct from the
    rank's ROW block (global column ids); the entries above the diagonal are redistributed to their column
    owners once, here.  Per-step buffers and index lists are fixed at construction.

    Attributes the step uses (all on the local device):
      xbuf = [x_own | x_recv]   own X rows, then the referenced rows of ranks 0..s-1 (owner order)
      xsend, send_idx           rows of x_own packed for ranks s+1.., and their local ids
      pull                      CSR  n_own x (n_own + n_xrecv): own row block, columns <= own block
      push                      CSR  n_psend x n_own: blocks (a, s), a < s, non-empty rows only, stacked by a
      psend / precv             partial output rows to ranks < s / from ranks > s
      add                       CSR  n_own x n_precv of ones: which received partial rows add into which row
    """

    def __init__(self, local_csr, bounds, rank, world, d, group=None, comm_stream=None, group_partials=None,
                 transport="nccl", push_ctas=0, hub_first=False, split_pull=False, fused_push=False, x_priority=0,
                 vblock_bounds=None, vblock_ids=None):
        """group_partials: a second communicator for the partial-row exchange, so that it can run while the
        X-row exchange is still in flight (opposite directions on full-duplex links); default: `group`.
        transport: "nccl" (all_to_all of packed rows) or "p2p" (gdr_push_rows_f32: rows are stored straight into
        the destination's receive region over NVLink; buffers live in symmetric memory).
        p2p only (both change the order of launches, not the arithmetic of any output row's partial sums):
          (any transport) vblock_bounds / vblock_ids: the cyclic virtual-block layout (cyclic_vblock_layout): bounds of
                      the virtual blocks in THIS numbering and their ids in the natural (hub-first) order.  The rule
                      "block (a, b) is computed by the owner of max(a, b)" is then applied to virtual blocks: an
                      entry (i, j) whose column lives on another rank is multiplied there (push) iff j's virtual
                      block comes after i's, else its X row is pulled.  Every pair of ranks exchanges in both
                      directions.
          fused_push  the push product stores its output rows STRAIGHT into the destination's receive region (the SpMM
                      kernel's row stores are peer stores over NVLink: 512 contiguous bytes per warp): one launch per
                      destination, rank 0's block first.  No psend buffer, no second copy kernel, and the partial rows
                      travel WHILE the product runs instead of after it — the partial exchange disappears from the
                      step's critical path and its copy kernel no longer competes with the pull product for SMs.
          x_priority  K > 0: ranks 0..K-1 (the hub ranks: they send the most X rows and are the last senders to finish)
                      push their X rows one destination at a time, HIGHEST rank first, and raise that destination's
                      arrival signal at once; every other rank keeps the single launch that serves all destinations
                      in proportion.  The last ranks (most rows, longest pull product) then hold all their X rows
                      when the second-largest sender is done instead of when rank 0 is.  Arrival is signalled per
                      source (put/wait_signal) instead of by one barrier.
          hub_first   the push product runs in two launches, the block whose partial rows go to rank 0 first, and
                      those rows leave as soon as it is done: rank 0 receives the most partial rows (its link is
                      the critical path of the partial exchange) and no longer waits for the whole push product.
          split_pull  the pull product is cut by SOURCE of the gathered X rows: "early" = own columns and rows of
                      ranks 1..s-1, "late" = rows of rank 0 (the hub block everybody references; its owner is the
                      last sender to finish).  Arrival is signalled per source (symmetric-memory put/wait_signal
                      instead of one barrier), the early product starts when ranks 1..s-1 have delivered, and the
                      late product is folded into the final add pass (same kernel, one pass over Y).  True: only
                      where an add pass exists; "force": also as an extra in-place pass (tests, 2-GPU runs)."""
        self.bounds = [int(b) for b in bounds]
        self.rank, self.world, self.d, self.group = rank, world, d, group
        self.group_partials = group_partials if group_partials is not None else group
        self.two_comms = group_partials is not None
        dev = local_csr.device
        s = rank
        lo, hi = self.bounds[s], self.bounds[s + 1]
        self.n, self.n_own = self.bounds[-1], hi - lo
        bnd = torch.tensor(self.bounds, dtype=torch.int64, device=dev)
        col = local_csr.col.to(torch.int64)
        deg = (local_csr.rowptr[1:] - local_csr.rowptr[:-1]).to(torch.int64)
        row = torch.repeat_interleave(torch.arange(self.n_own, device=dev), deg)
        owner = torch.searchsorted(bnd, col, right=True) - 1
        val = local_csr.val

        # ---- pull part: columns owned by ranks <= s (virtual-block layout: own columns and earlier blocks) ---------
        self.cyclic = vblock_bounds is not None
        if self.cyclic:
            vbnd = torch.as_tensor(np.asarray(vblock_bounds), dtype=torch.int64, device=dev)
            vid = torch.as_tensor(np.asarray(vblock_ids), dtype=torch.int64, device=dev)
            nvb = vid.numel()
            k_col = vid[torch.clamp(torch.searchsorted(vbnd, col, right=True) - 1, 0, nvb - 1)]
            k_row = vid[torch.clamp(torch.searchsorted(vbnd, row + lo, right=True) - 1, 0, nvb - 1)]
            keep = (owner == s) | (k_col < k_row)
            del k_col, k_row
        else:
            keep = owner <= s
        kcol, krow, kval = col[keep], row[keep], val[keep]
        need_from, new_col, _ = plan_exchange(kcol, self.bounds, s)       # owners > s get empty lists
        self.xrecv_splits = [int(t.numel()) for t in need_from]
        n_xrecv = int(sum(self.xrecv_splits))
        self.pull = csr_from_sorted_rows(krow, new_col, kval, self.n_own, self.n_own + n_xrecv)
        # entries that gather rows received from rank 0 (they sit first in the receive region)
        self._late_mask = (new_col >= self.n_own) & (new_col < self.n_own + self.xrecv_splits[0]) if s > 0 else None
        self._pull_parts = (krow, new_col, kval)
        want = torch.cat(need_from) if n_xrecv else torch.zeros(0, dtype=torch.int64, device=dev)
        self.send_idx, self.xsend_splits = _a2a_var(want, self.xrecv_splits, group)

        # ---- push part: entries above the diagonal go to their column owner, once ----------------------
        up = ~keep
        uo = owner[up]
        order = torch.argsort(uo, stable=True)                            # by owner, row-major inside
        cnt = torch.bincount(uo, minlength=world).tolist()
        g_row, _ = _a2a_var((row[up] + lo)[order], cnt, group)
        g_col, _ = _a2a_var(col[up][order], cnt, group)
        g_val, from_counts = _a2a_var(val[up][order], cnt, group)         # arrives grouped by source rank
        src = torch.repeat_interleave(torch.arange(world, device=dev),
                                      torch.tensor(from_counts, dtype=torch.int64, device=dev))
        # non-empty rows per source block, in (source, row) order = the layout of psend
        rows_u, inv = torch.unique(g_row, return_inverse=True)            # sorted global rows => grouped by source
        self.n_psend = int(rows_u.numel())
        self.push = csr_from_sorted_rows(inv, g_col - lo, g_val, self.n_psend, self.n_own)
        cut = torch.searchsorted(rows_u, bnd)
        self.psend_splits = [int(cut[a + 1] - cut[a]) for a in range(world)]
        back = rows_u - bnd[torch.searchsorted(bnd, rows_u, right=True) - 1]   # owner-local row ids
        add_rows, self.precv_splits = _a2a_var(back, self.psend_splits, group)
        self.n_precv = int(add_rows.numel())
        pos = torch.arange(self.n_precv, device=dev)
        o2 = torch.argsort(add_rows, stable=True)                         # by row; sources stay in rank order
        self.add = csr_from_sorted_rows(add_rows[o2], pos[o2], torch.ones(self.n_precv, dtype=val.dtype, device=dev),
                                        self.n_own, max(self.n_precv, 1))
        self._add_parts = (add_rows[o2], pos[o2])
        del src
        # rows of the push product whose partial rows go to rank 0 come first (psend is stacked by destination)
        self.n_psend_hub = self.psend_splits[0] if s > 0 else 0

        # ---- per-step buffers ---------------------------------------------------------------------------
        f = dict(dtype=val.dtype, device=dev)
        self.p2p = transport == "p2p"
        self.push_ctas = int(push_ctas)
        self.x_priority = int(x_priority) if (transport == "p2p" and world > 1 and vblock_bounds is None) else 0
        self.fused_push = bool(fused_push) and self.p2p and world > 1
        self.hub_first = bool(hub_first) and self.p2p and world > 2 and not self.fused_push and not self.cyclic
        self.split_pull = split_pull if (self.p2p and world > 1 and not self.cyclic) else False
        self._split_requested = self.split_pull                           # same value on every rank
        self._push_slices = {}
        if self.p2p:
            self._setup_p2p(n_xrecv, f)
            self._setup_split(val)
        else:
            self.xbuf = torch.empty(self.n_own + n_xrecv, d, **f)
            self.precv = torch.empty(max(self.n_precv, 1), d, **f)[:self.n_precv]
            self.xsend = torch.empty(int(self.send_idx.numel()), d, **f)
        self._pull_parts = self._add_parts = self._late_mask = None        # construction-time index lists
        self.x_own, self.x_recv = self.xbuf[:self.n_own], self.xbuf[self.n_own:]
        self.psend = torch.empty(0 if self.fused_push else self.n_psend, d, **f)
        self.y = torch.empty(self.n_own, d, **f)
        self.recv_rows = n_xrecv + self.n_precv
        self.sent_rows = int(self.send_idx.numel()) + self.n_psend
        self.nnz_local = self.pull.nnz + self.push.nnz
        self.comm_stream = comm_stream
        if self.comm_stream is None and dev.type == "cuda":
            self.comm_stream = torch.cuda.Stream(device=dev, priority=-1)
        self.comm_stream2 = (torch.cuda.Stream(device=dev, priority=-1)
                             if ((self.two_comms or self.p2p) and dev.type == "cuda") else self.comm_stream)
        self.transport = ("p2p: X rows by gdr_push_rows_f32, partial rows stored into peer memory by the push product itself"
                          if self.fused_push else
                          "p2p stores into peer symmetric memory (gdr_push_rows_f32) x2" if self.p2p else
                          "nccl all_to_all x2") + " (X rows to higher ranks, partial Y rows to lower ranks), L-shaped partition"

    def _setup_p2p(self, n_xrecv, f):
        """xbuf and precv of every rank in ONE symmetric allocation; destination addresses of both pushes."""
        import ctypes
        import torch.distributed._symmetric_memory as symm
        s, world, d = self.rank, self.world, self.d
        meta = [None] * world
        dist.all_gather_object(meta, dict(n_own=self.n_own, xrecv=self.xrecv_splits, precv=self.precv_splits),
                               group=self.group)
        xrows_max = max(m["n_own"] + sum(m["xrecv"]) for m in meta)
        prows_max = max(max(sum(m["precv"]) for m in meta), 1)
        self._symm_t = symm.empty((xrows_max + prows_max) * d, dtype=f["dtype"], device=f["device"])
        gname = (self.group or dist.group.WORLD).group_name
        self._hdl = symm.rendezvous(self._symm_t, gname)
        buf = self._symm_t.view(xrows_max + prows_max, d)
        self.xbuf = buf[:self.n_own + n_xrecv]
        self.precv = buf[xrows_max:xrows_max + self.n_precv]
        self.xsend = torch.empty(0, d, **f)                                  # no pack buffer on this transport
        esz = self._symm_t.element_size()
        base = [int(p) for p in self._hdl.buffer_ptrs]
        arr, i64 = ctypes.c_void_p * world, ctypes.c_int64 * world
        # X rows: to every r > s, into r's xbuf after r's own rows and the rows of ranks < s
        xdst, xidx, xcnt, off = [None] * world, [None] * world, [0] * world, 0
        for r in range(world):
            c = self.xsend_splits[r]
            if c:
                xdst[r] = base[r] + (meta[r]["n_own"] + sum(meta[r]["xrecv"][:s])) * d * esz
                xidx[r] = self.send_idx.data_ptr() + off * 8
                xcnt[r] = c
            off += c
        self._xpush = (arr(*xdst), arr(*xidx), i64(*([0] * world)), i64(*xcnt))
        self._xpush_one = {}
        for r in range(world):
            if xcnt[r]:
                one = [0] * world
                one[r] = xcnt[r]
                self._xpush_one[r] = (arr(*xdst), arr(*xidx), i64(*([0] * world)), i64(*one))
        # partial rows: to every a < s, contiguous slices of psend, into a's precv after the rows of ranks < s
        pdst, pcnt, poff, off = [None] * world, [0] * world, [0] * world, 0
        for a in range(world):
            c = self.psend_splits[a]
            if c:
                pdst[a] = base[a] + (xrows_max + sum(meta[a]["precv"][:s])) * d * esz
                poff[a], pcnt[a] = off, c
            off += c
        self._ppush = (arr(*pdst), arr(*([None] * world)), i64(*poff), i64(*pcnt))
        # hub-first variant: destination 0 alone, then the rest
        hub_cnt = [pcnt[0]] + [0] * (world - 1)
        rest_cnt = [0] + pcnt[1:]
        self._ppush_hub = (arr(*pdst), arr(*([None] * world)), i64(*poff), i64(*hub_cnt))
        self._ppush_rest = (arr(*pdst), arr(*([None] * world)), i64(*poff), i64(*rest_cnt))
        # fused push: the destination's receive slice as a local tensor over peer memory
        self._push_ranges = []
        if self.fused_push:
            for a in range(world):
                if pcnt[a]:
                    view = self._hdl.get_buffer(a, (pcnt[a], d), f["dtype"], (xrows_max + sum(meta[a]["precv"][:s])) * d)
                    self._push_ranges.append((poff[a], poff[a] + pcnt[a], view))
        self._xrows_max = xrows_max
        self._x_sources = [r for r in range(world) if self.xrecv_splits[r] > 0]
        self._x_dests = [r for r in range(world) if self.xsend_splits[r] > 0]

    def _setup_split(self, val):
        """split_pull: pull_early (own columns + rows of ranks 1..s-1), and `final` = [late pull entries | partial-row
        adds] as ONE CSR over the symmetric buffer (xbuf and precv live in the same allocation), so that the late
        product costs no extra pass over Y.  Entry order inside a row: late pull entries by ascending column, then
        the partial rows by source rank - fixed, so the result is deterministic."""
        self.pull_early = self.pull_late = self.final = None
        if not self.split_pull or self.rank == 0 or self.xrecv_splits[0] == 0:
            self.split_pull = False
            return
        if self.split_pull != "force" and self.n_precv == 0:
            self.split_pull = False                       # no add pass to fold the late product into
            return
        krow, new_col, kval = self._pull_parts
        late = self._late_mask
        dev = krow.device
        n_cols = self.n_own + int(sum(self.xrecv_splits))
        self.pull_early = csr_from_sorted_rows(krow[~late], new_col[~late], kval[~late], self.n_own, n_cols)
        if self.n_precv:
            add_rows, add_pos = self._add_parts
            rows = torch.cat([krow[late], add_rows])
            cols = torch.cat([new_col[late], add_pos + self._xrows_max])
            vals = torch.cat([kval[late], torch.ones(self.n_precv, dtype=val.dtype, device=dev)])
            order = torch.argsort(rows, stable=True)      # late entries stay ahead of the adds inside every row
            self.final = csr_from_sorted_rows(rows[order], cols[order], vals[order], self.n_own,
                                              self._xrows_max + self.n_precv)
            self._buf_all = self._symm_t.view(-1, self.d)[:self._xrows_max + self.n_precv]
        else:
            self.pull_late = csr_from_sorted_rows(krow[late], new_col[late], kval[late], self.n_own, n_cols)

    def _push(self, src, table):
        dst, idx, off, cnt = table
        _ffi.check(_ffi.lib().gdr_push_rows_f32(_ffi.ptr(src) if src.numel() else None, self.d, self.world, dst, idx,
                                                off, cnt, self.push_ctas, _ffi.stream()), "push_rows")

    @classmethod
    def from_full(cls, full_csr, bounds, rank, world, d, group=None, group_partials=None, **kw):
        lo, hi = int(bounds[rank]), int(bounds[rank + 1])
        p0, p1 = int(full_csr.rowptr[lo].item()), int(full_csr.rowptr[hi].item())
        local = DeviceCSR((full_csr.rowptr[lo:hi + 1] - p0).contiguous(), full_csr.col[p0:p1].clone(),
                          full_csr.val[p0:p1].clone(), hi - lo, full_csr.n_cols)
        return cls(local, bounds, rank, world, d, group, group_partials=group_partials, **kw)

    @classmethod
    def from_full_cyclic(cls, full_csr, vb, rank, world, d, group=None, group_partials=None, **kw):
        """Partition of `full_csr` (natural vertex order) under the cyclic virtual-block layout of `vb`
        (cyclic_vblock_layout).  The object works in the renumbered order: `rows_old` holds the natural ids of this
        rank's rows — x_own = X[rows_old], and spmm() returns (A X)[rows_old]."""
        dev = full_csr.device
        order, new_vb, rank_bounds = cyclic_vblock_layout(vb, world)
        perm, inv = cyclic_permutation(vb, order, dev)
        lo, hi = int(rank_bounds[rank]), int(rank_bounds[rank + 1])
        rows_old = perm[lo:hi]
        rp = full_csr.rowptr.to(torch.int64)
        deg = rp[rows_old + 1] - rp[rows_old]
        rowptr = torch.zeros(hi - lo + 1, dtype=torch.int64, device=dev)
        rowptr[1:] = torch.cumsum(deg, 0)
        nnz = int(rowptr[-1].item())
        # entries of the gathered rows: position = start of the old row + offset inside the row
        row_new = torch.repeat_interleave(torch.arange(hi - lo, device=dev), deg)
        src = rp[rows_old][row_new] + (torch.arange(nnz, device=dev) - rowptr[:-1][row_new])
        col = inv[full_csr.col[src].to(torch.int64)]
        val = full_csr.val[src]
        del src
        key = row_new * int(rank_bounds[-1]) + col              # ascending new column ids inside every row
        o = torch.argsort(key)
        del key, row_new
        local = DeviceCSR(rowptr.to(torch.int32), col[o].to(torch.int32).contiguous(), val[o].contiguous(), hi - lo,
                          full_csr.n_cols)
        del col, val, o
        part = cls(local, rank_bounds, rank, world, d, group, group_partials=group_partials, vblock_bounds=new_vb,
                   vblock_ids=order, **kw)
        part.rows_old = rows_old
        return part

    # the two exchanges, separately callable (tests replace the device products around them)
    SIG_X = 4                                       # signal-pad channel of the per-source X arrival flags

    def exchange_x(self):
        """Returns (ev_early, ev_late): events on the current stream after which the rows of ranks 1..s-1 / of every
        source have landed in xbuf.  Without split_pull both are the same event (one barrier)."""
        if self.p2p:
            cur = torch.cuda.current_stream()
            self._hdl.barrier(channel=0)            # every rank is done reading last step's xbuf / precv
            self._ev_b0 = cur.record_event()
            if self.x_priority:
                if self.rank < self.x_priority:
                    for r in sorted(self._x_dests, reverse=True):      # one destination at a time, highest first
                        self._push(self.x_own, self._xpush_one[r])
                        self._hdl.put_signal(r, channel=self.SIG_X)
                else:
                    if self.send_idx.numel():
                        self._push(self.x_own, self._xpush)
                    for r in self._x_dests:
                        self._hdl.put_signal(r, channel=self.SIG_X)
                for r in self._x_sources:
                    self._hdl.wait_signal(r, channel=self.SIG_X)
                ev = cur.record_event()
                return ev, ev
            if self.send_idx.numel():
                self._push(self.x_own, self._xpush)
            if self._signals:
                for r in self._x_dests:             # my rows for r have landed (kernel completion orders the stores)
                    self._hdl.put_signal(r, channel=self.SIG_X)
                for r in self._x_sources:
                    if r != 0:
                        self._hdl.wait_signal(r, channel=self.SIG_X)
                ev_early = cur.record_event()
                if 0 in self._x_sources:
                    self._hdl.wait_signal(0, channel=self.SIG_X)
                return ev_early, cur.record_event()
            self._hdl.barrier(channel=1)            # every rank's rows have landed
            ev = cur.record_event()
            return ev, ev
        if self.xsend.numel():
            _ffi.check(_ffi.lib().gdr_gather_rows_f32(_ffi.ptr(self.x_own), _ffi.ptr(self.send_idx),
                                                      self.send_idx.numel(), self.n_own, self.d,
                                                      _ffi.ptr(self.xsend), _ffi.stream()), "gather_rows")
        dist.all_to_all_single(self.x_recv, self.xsend, self.xrecv_splits, self.xsend_splits, group=self.group)
        ev = torch.cuda.current_stream().record_event() if self.x_own.is_cuda else None
        return ev, ev

    @property
    def _signals(self):
        """Per-source arrival signals replace barrier 1 when ANY rank splits its pull product (collective choice:
        the option is the same on every rank, and a sender cannot know which receivers fold their late part)."""
        return bool(self._split_requested)

    def push_product(self):
        if self.n_psend:
            self.push.spmm(self.x_own, out=self.psend)                    # needs only the own X block

    def _push_product_rows(self, lo, hi, out=None):
        """Rows [lo, hi) of the push product (a row slice of the same CSR: same kernel, same per-row arithmetic);
        out: where the rows go (default: psend[lo:hi]; fused_push: the destination's receive slice, peer memory)."""
        if hi <= lo:
            return
        key = (lo, hi)
        sub = self._push_slices.get(key)
        if sub is None:
            rp = self.push.rowptr
            p0, p1 = int(rp[lo].item()), int(rp[hi].item())
            sub = DeviceCSR((rp[lo:hi + 1] - p0).contiguous(), self.push.col[p0:p1], self.push.val[p0:p1], hi - lo,
                            self.push.n_cols)
            self._push_slices[key] = sub
        sub.spmm(self.x_own, out=self.psend[lo:hi] if out is None else out)

    def hub_reduce(self):
        pass                                                              # (NvlsLShapedSpMM only)

    def exchange_partials(self, table=None, last=True):
        if self.p2p:
            torch.cuda.current_stream().wait_event(self._ev_b0)
            if self.n_psend:
                self._push(self.psend, table if table is not None else self._ppush)
            if last:
                self._hdl.barrier(channel=2)
            return
        dist.all_to_all_single(self.precv, self.psend, self.precv_splits, self.psend_splits,
                               group=self.group_partials)

    def spmm(self, x_local=None):
        """Y_block = (Â·X)[lo:hi]; x_local = this rank's rows of X (None: already written into self.x_own)."""
        if x_local is not None and x_local.data_ptr() != self.x_own.data_ptr():
            self.x_own.copy_(x_local)
        cur = torch.cuda.current_stream()
        cs, cs2 = self.comm_stream, self.comm_stream2
        cs.wait_stream(cur)
        with torch.cuda.stream(cs):
            ev_early, ev_late = self.exchange_x()
        if self.fused_push:
            cur.wait_event(self._ev_b0)                                   # every rank is done reading last step's precv
            for lo_, hi_, view in self._push_ranges:                      # rank 0's block first: it receives the most
                self._push_product_rows(lo_, hi_, out=view)
            ev_p = cur.record_event()
            with torch.cuda.stream(cs2):
                cs2.wait_event(ev_p)
                self._hdl.barrier(channel=2)                              # every rank's partial rows have landed
                ev_y = cs2.record_event()
        elif self.hub_first and self.n_psend_hub and self.n_psend_hub < self.n_psend:
            self._push_product_rows(0, self.n_psend_hub)                  # partial rows for rank 0 first ...
            ev_p0 = cur.record_event()
            with torch.cuda.stream(cs2):
                cs2.wait_event(ev_p0)
                self.exchange_partials(self._ppush_hub, last=False)       # ... and on their way at once
            self._push_product_rows(self.n_psend_hub, self.n_psend)
            ev_p = cur.record_event()
            with torch.cuda.stream(cs2):
                cs2.wait_event(ev_p)
                self.exchange_partials(self._ppush_rest, last=True)
                ev_y = cs2.record_event()
        else:
            self.push_product()
            ev_p = cur.record_event()
            with torch.cuda.stream(cs2):
                cs2.wait_event(ev_p)
                self.exchange_partials()
                ev_y = cs2.record_event()
        if self.split_pull:
            cur.wait_event(ev_early)
            self.pull_early.spmm(self.xbuf, out=self.y)                   # own columns + rows of ranks 1..s-1
            cur.wait_event(ev_late)
            cur.wait_event(ev_y)
            if self.final is not None:
                self.final.spmm(self._buf_all, e1=self.y, out=self.y)     # rank 0's rows + the partial rows, one pass
            else:
                self.pull_late.spmm(self.xbuf, e1=self.y, out=self.y)
            return self.y
        if self.x_recv.shape[0]:
            cur.wait_event(ev_late)
        self.pull.spmm(self.xbuf, out=self.y)
        if ev_late is not None:
            # order the caller's next write into x_own after the exchange that READS it (rank 0 pulls nothing, yet
            # its rows are still being pushed while its pull product runs); off the critical path here
            cur.wait_event(ev_late)
        cur.wait_event(ev_y)
        if self.n_precv:
            self.add.spmm(self.precv, e1=self.y, out=self.y)              # fixed (row, source rank) order
        return self.y


# ---- a timeline model of one L-shaped step, used to refine the bounds ------------------------------------------
# Constants measured on B200 / NVLink 5 (profiles/diag_r1_n8_lshape_hubs0.txt): seconds per gathered entry, per
# output row written by an SpMM, per row touched by the add pass, and per 512-byte row moved over one NVSwitch port.
# volume_scale: measured distinct-row counts over the Poisson estimate on R-MAT scale 24 (0.45-0.6 for every rank
# pair: inside a histogram bin the degrees are far from uniform, which the estimate ignores).
STEP_MODEL = dict(t_nnz=50e-12, t_row=250e-12, t_add=235e-12, link_bytes_per_s=695e9, t_sync=0.15e-3,
                  volume_scale=0.5)


def lshape_step_model(hist, bin_size, n, bbins, d, consts=None):
    """Predicted step time [s] and per-rank detail for L-shape bounds given in bins (bbins[0] = 0, bbins[-1] = B).
    Block (r, s) is multiplied by rank max(r, s); volumes are Poisson estimates from the coarse histogram."""
    k = dict(STEP_MODEL, **(consts or {}))
    hist = np.asarray(hist, dtype=np.float64)
    nb = hist.shape[0]
    p = len(bbins) - 1
    rows_in = np.maximum(np.minimum(np.arange(1, nb + 1) * bin_size, n) - np.minimum(np.arange(nb) * bin_size, n), 1.0)
    owner = np.searchsorted(np.asarray(bbins), np.arange(nb), side="right") - 1        # bin -> rank
    own = np.array([rows_in[bbins[r]:bbins[r + 1]].sum() for r in range(p)])
    nnz = np.zeros((p, p))
    pullr = np.zeros((p, p))      # [r, s] r > s: distinct X rows of s that r references
    pushr = np.zeros((p, p))      # [r, s] r < s: non-empty rows of block (r, s)
    for r in range(p):
        strip = hist[bbins[r]:bbins[r + 1], :]                                          # rows of r
        col_tot = strip.sum(0)
        hit_c = rows_in * (1.0 - np.exp(-col_tot / rows_in))                            # distinct columns per col bin
        np.add.at(nnz[r], owner, col_tot)
        np.add.at(pullr[r], owner, hit_c)
        rr = rows_in[bbins[r]:bbins[r + 1]]
        for s in range(r + 1, p):
            lam = strip[:, bbins[s]:bbins[s + 1]].sum(1) / rr
            pushr[r, s] = (rr * (1.0 - np.exp(-lam))).sum()
    lower = np.tril(np.ones((p, p)), -1)
    upper = lower.T
    nnz_pull = (nnz * (lower + np.eye(p))).sum(1)
    nnz_push = (nnz * upper).sum(0)
    vs = k["volume_scale"]
    recv_x, sent_x = vs * (pullr * lower).sum(1), vs * (pullr * lower).sum(0)
    recv_p, sent_p = vs * (pushr * upper).sum(1), vs * (pushr * upper).sum(0)
    t_push = k["t_nnz"] * nnz_push + k["t_row"] * sent_p
    t_pull = k["t_nnz"] * nnz_pull + k["t_row"] * own
    t_add = np.where(recv_p > 0, k["t_add"] * (own + recv_p), 0.0)
    rowb = d * 4.0
    T_x = max(sent_x.max(), recv_x.max()) * rowb / k["link_bytes_per_s"] + k["t_sync"]
    T_p = max(sent_p.max(), recv_p.max()) * rowb / k["link_bytes_per_s"] + k["t_sync"]
    pull_start = np.where(recv_x > 0, np.maximum(t_push, T_x), t_push)
    part_done = t_push.max() + T_p                        # one barrier: the slowest sender gates everybody
    finish = np.maximum(pull_start + t_pull, np.where(recv_p > 0, part_done, 0.0)) + t_add
    return float(finish.max()), dict(finish=finish, t_push=t_push, t_pull=t_pull, t_add=t_add, T_x=T_x, T_p=T_p,
                                     sent_x=sent_x, recv_x=recv_x, sent_p=sent_p, recv_p=recv_p, own=own,
                                     nnz_pull=nnz_pull, nnz_push=nnz_push)


def refine_lshape_bounds(hist, bin_size, n, bounds, d, consts=None, sweeps=6):
    """Coordinate descent on the interior bounds (in histogram bins) against lshape_step_model, starting from
    `bounds` (rows).  Deterministic; every rank computes the same result from the same histogram."""
    nb = np.asarray(hist).shape[0]
    bb = [int(round(b / bin_size)) for b in bounds]
    bb[0], bb[-1] = 0, nb
    best, _ = lshape_step_model(hist, bin_size, n, bb, d, consts)
    for _ in range(sweeps):
        improved = False
        for i in range(1, len(bb) - 1):
            for step in (64, 16, 4, 1):
                moved = True
                while moved:
                    moved = False
                    for cand in (bb[i] - step, bb[i] + step):
                        if bb[i - 1] < cand < bb[i + 1]:
                            trial = bb[:i] + [cand] + bb[i + 1:]
                            t, _ = lshape_step_model(hist, bin_size, n, trial, d, consts)
                            if t < best * (1.0 - 1e-4):
                                best, bb, moved, improved = t, trial, True, True
                                break
        if not improved:
            break
    return np.minimum(np.asarray(bb, dtype=np.int64) * bin_size, n), best
