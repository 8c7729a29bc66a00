"""Row-partitioned SpMM for graphs that exceed one GPU (north_star: R-MAT 16M nodes / 512M edges over the
8 B200s of one box).  One process per GPU (torch.distributed, NCCL over NVLink 5 / NVSwitch).

Partition: p contiguous row blocks of Â balanced by nnz (R-MAT is skewed); rank r owns rows [lo_r, hi_r) of
Â, of X and of Y.  Every rank needs all of X for its gathers, so a step is
    all-gather of the X row blocks (each rank contributes N*d*s/p bytes)  ->  local CSR SpMM
and the output rows stay local (they are the next layer's X block).  Âᵀ·G (backward) uses a second
partition of Âᵀ built the same way, so accumulation order stays deterministic (no reduce-scatter).

The reference has no distributed code at all; this module is new surface, kept to the one
exchange step the path really has.
"""
import numpy as np
import torch
import torch.distributed as dist

from .graph import DeviceCSR


def nnz_balanced_bounds(rowptr, parts):
    """Row boundaries [parts+1] of contiguous blocks with ~equal nnz.  rowptr: numpy int array [n+1]."""
    rowptr = np.asarray(rowptr)
    nnz = int(rowptr[-1])
    targets = (np.arange(1, parts) * (nnz / parts)).astype(np.int64)
    cuts = np.searchsorted(rowptr, targets, side="left")
    b = np.concatenate([[0], cuts, [len(rowptr) - 1]]).astype(np.int64)
    return np.maximum.accumulate(b)


class RowPartitionedSpMM:
    """This rank's row block of Â plus the buffers of the exchange step."""

    def __init__(self, local_csr, bounds, rank, world, d, group=None):
        self.csr = local_csr                   # rows [lo, hi), columns global
        self.bounds = [int(b) for b in bounds]
        self.rank, self.world, self.d = rank, world, d
        self.group = group
        self.n = self.bounds[-1]
        dev = local_csr.device
        self.x_full = torch.empty(self.n, d, dtype=torch.float32, device=dev)
        self.views = [self.x_full[self.bounds[r]:self.bounds[r + 1]] for r in range(world)]
        self.y = torch.empty(local_csr.n_rows, d, dtype=torch.float32, device=dev)
        self.transport = "nccl all_gather (uneven row blocks) then local SpMM"

    @classmethod
    def from_full(cls, full_csr, bounds, rank, world, d, group=None):
        lo, hi = int(bounds[rank]), int(bounds[rank + 1])
        p0, p1 = int(full_csr.rowptr[lo].item()), int(full_csr.rowptr[hi].item())
        local = DeviceCSR((full_csr.rowptr[lo:hi + 1] - p0).contiguous(), full_csr.col[p0:p1].clone(),
                          full_csr.val[p0:p1].clone(), hi - lo, full_csr.n_cols)
        return cls(local, bounds, rank, world, d, group)

    def spmm(self, x_local):
        """x_local: this rank's rows of X [hi-lo, d].  Returns this rank's rows of Y = Â·X."""
        if self.world == 1:
            return self.csr.spmm(x_local, out=self.y)
        dist.all_gather(self.views, x_local.contiguous(), group=self.group)
        return self.csr.spmm(self.x_full, out=self.y)
