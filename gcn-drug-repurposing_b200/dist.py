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
