"""DeviceCSR — the adjacency as the kernels want it: int32 rowptr/col + values in HBM, plus a
lazily built transposed copy for the backward pass (Âᵀ·G).

The reference hands the model an *uncoalesced int64 COO* torch sparse tensor
(helpers/helper.py:92-96, consumed by torch.sparse.mm at modules/model.py:163,169).  The
drop-in layer converts it once on the device (gdr_csr_from_coo) and caches the result on the
adjacency tensor object, which train.py passes unchanged every iteration (train.py:161).
The transpose is always explicit: Â is not assumed symmetric.
"""
import ctypes

import numpy as np
import torch

from . import _ffi


class DeviceCSR:
    # benchmark hook: when set to a list, every spmm() appends
    # (start_event, end_event, nnz, n_rows, d, has_mul, has_add) recorded on the launching stream
    timer = None

    def __init__(self, rowptr, col, val, n_rows, n_cols):
        self.rowptr = rowptr  # int32 [n_rows+1]
        self.col = col        # int32 [nnz]
        self.val = val        # float32 / float64 [nnz]
        self.n_rows = int(n_rows)
        self.n_cols = int(n_cols)
        self._t = None
        self._plan = None

    # ---- constructors -------------------------------------------------------------------
    @classmethod
    def from_coo(cls, row, col, val, n_rows, n_cols):
        """row/col: int64 CUDA tensors (any order, duplicates kept); val: fp32/fp64 CUDA tensor."""
        L = _ffi.lib()
        dev = row.device
        nnz = int(row.numel())
        row = row.contiguous()
        col = col.contiguous()
        val = val.contiguous()
        rowptr = torch.empty(n_rows + 1, dtype=torch.int32, device=dev)
        colo = torch.empty(max(nnz, 1), dtype=torch.int32, device=dev)[:nnz]
        perm = torch.empty(max(nnz, 1), dtype=torch.int64, device=dev)[:nnz]
        ws_bytes = L.gdr_csr_from_coo_workspace_bytes(nnz, n_rows, n_cols)
        ws = _ffi.workspace(ws_bytes, dev)
        _ffi.check(L.gdr_csr_from_coo(_ffi.ptr(row), _ffi.ptr(col), nnz, n_rows, n_cols, _ffi.ptr(rowptr),
                                      _ffi.ptr(colo), _ffi.ptr(perm), _ffi.ptr(ws), ws.numel(), _ffi.stream()),
                   "csr_from_coo")
        valo = torch.empty_like(val)
        _ffi.check(L.gdr_permute_values(_ffi.ptr(val), _ffi.ptr(perm), nnz, val.element_size(), _ffi.ptr(valo),
                                        _ffi.stream()), "permute_values")
        return cls(rowptr, colo, valo, n_rows, n_cols)

    @classmethod
    def from_torch_sparse(cls, adj):
        """adj: torch sparse COO tensor on a CUDA device (what convert_sparse_matrix_to_sparse_tensor
        returns, moved with .cuda() as train.py:120 does)."""
        if not adj.is_cuda:
            raise RuntimeError("DeviceCSR: the adjacency must live on a CUDA device (no CPU path)")
        if adj.layout != torch.sparse_coo:
            raise RuntimeError(f"DeviceCSR: expected a sparse COO adjacency, got {adj.layout}")
        idx = adj._indices()
        return cls.from_coo(idx[0], idx[1], adj._values(), adj.shape[0], adj.shape[1])

    @classmethod
    def from_scipy(cls, m, device, dtype=torch.float32):
        """m: scipy csr_matrix (indices need not be sorted: goes through the COO path)."""
        m = m.tocsr()
        if m.has_sorted_indices and m.has_canonical_format:
            rowptr = torch.from_numpy(np.asarray(m.indptr, dtype=np.int32)).to(device)
            col = torch.from_numpy(np.asarray(m.indices, dtype=np.int32)).to(device)
            val = torch.from_numpy(np.asarray(m.data)).to(device=device, dtype=dtype)
            return cls(rowptr, col, val, m.shape[0], m.shape[1])
        coo = m.tocoo()
        return cls.from_coo(torch.from_numpy(coo.row.astype(np.int64)).to(device),
                            torch.from_numpy(coo.col.astype(np.int64)).to(device),
                            torch.from_numpy(coo.data).to(device=device, dtype=dtype), m.shape[0], m.shape[1])

    # ---- properties ------------------------------------------------------------------------
    @property
    def nnz(self):
        return int(self.col.numel())

    @property
    def device(self):
        return self.rowptr.device

    def to_scipy(self):
        import scipy.sparse as sp
        return sp.csr_matrix((self.val.cpu().numpy(), self.col.cpu().numpy(), self.rowptr.cpu().numpy()),
                             shape=(self.n_rows, self.n_cols))

    # ---- transpose ----------------------------------------------------------------------------
    def transpose(self):
        if self._t is None:
            self.transpose_with_perm()
        return self._t

    def transpose_with_perm(self):
        """(transposed DeviceCSR, perm int64[nnz]) with perm[i] = position in self of transposed entry i."""
        L = _ffi.lib()
        dev = self.device
        nnz = self.nnz
        t_rowptr = torch.empty(self.n_cols + 1, dtype=torch.int32, device=dev)
        t_col = torch.empty(max(nnz, 1), dtype=torch.int32, device=dev)[:nnz]
        perm = torch.empty(max(nnz, 1), dtype=torch.int64, device=dev)[:nnz]
        ws = _ffi.workspace(L.gdr_csr_transpose_workspace_bytes(nnz, self.n_rows, self.n_cols), dev)
        _ffi.check(L.gdr_csr_transpose(_ffi.ptr(self.rowptr), _ffi.ptr(self.col), nnz, self.n_rows, self.n_cols,
                                       _ffi.ptr(t_rowptr), _ffi.ptr(t_col), _ffi.ptr(perm), _ffi.ptr(ws),
                                       ws.numel(), _ffi.stream()), "csr_transpose")
        t_val = torch.empty_like(self.val)
        _ffi.check(L.gdr_permute_values(_ffi.ptr(self.val), _ffi.ptr(perm), nnz, self.val.element_size(),
                                        _ffi.ptr(t_val), _ffi.stream()), "permute_values")
        self._t = DeviceCSR(t_rowptr, t_col, t_val, self.n_cols, self.n_rows)
        self._t._t = self
        return self._t, perm

    # ---- (A + I) symmetric normalisation: preprocess_graph on the device -----------------------
    def normalized_with_identity(self):
        """D^-1/2 (A+I) D^-1/2 in fp64 (helpers/helper.py:82-89).  Returns a new fp64 DeviceCSR."""
        if self.n_rows != self.n_cols:
            raise RuntimeError("preprocess_graph needs a square adjacency")
        if self.val.dtype != torch.float64:
            raise RuntimeError("preprocess_graph runs in fp64 like the reference (got %s)" % self.val.dtype)
        L = _ffi.lib()
        dev = self.device
        n = self.n_rows
        out_rowptr = torch.empty(n + 1, dtype=torch.int32, device=dev)
        ws = _ffi.workspace(L.gdr_csr_add_identity_workspace_bytes(n), dev)
        nnz_out = ctypes.c_int64(0)
        _ffi.check(L.gdr_csr_add_identity_count(_ffi.ptr(self.rowptr), _ffi.ptr(self.col), n, _ffi.ptr(out_rowptr),
                                                ctypes.byref(nnz_out), _ffi.ptr(ws), ws.numel(), _ffi.stream()),
                   "csr_add_identity_count")
        nnz = int(nnz_out.value)
        out_col = torch.empty(max(nnz, 1), dtype=torch.int32, device=dev)[:nnz]
        out_val = torch.empty(max(nnz, 1), dtype=torch.float64, device=dev)[:nnz]
        dis = torch.empty(max(n, 1), dtype=torch.float64, device=dev)[:n]
        _ffi.check(L.gdr_csr_normalize_f64(_ffi.ptr(self.rowptr), _ffi.ptr(self.col), _ffi.ptr(self.val), n,
                                           _ffi.ptr(out_rowptr), _ffi.ptr(out_col), _ffi.ptr(out_val),
                                           _ffi.ptr(dis), _ffi.stream()), "csr_normalize")
        out = DeviceCSR(out_rowptr, out_col, out_val, n, n)
        out.deg_inv_sqrt = dis
        return out

    def astype(self, dtype):
        if self.val.dtype == dtype:
            return self
        return DeviceCSR(self.rowptr, self.col, self.val.to(dtype), self.n_rows, self.n_cols)

    # ---- degree-aware schedule --------------------------------------------------------------------
    HUB_THRESHOLD = 512    # rows longer than this are cut into chunks ...
    HUB_CHUNK = 256        # ... of this many entries, one warp each

    def plan(self):
        """Hub-row schedule for gdr_spmm_csr_planned_f32, built once per matrix (index arithmetic on the row
        pointers; see the SpmmPlan comment in csrc/spmm.cu).  None when no row exceeds HUB_THRESHOLD."""
        if self._plan is None:
            dev = self.device
            deg = (self.rowptr[1:] - self.rowptr[:-1]).to(torch.int64)
            hubs = torch.nonzero(deg > self.HUB_THRESHOLD).flatten()
            if hubs.numel() == 0:
                self._plan = False
            else:
                nch = (deg[hubs] + self.HUB_CHUNK - 1) // self.HUB_CHUNK
                first = torch.cumsum(nch, 0) - nch
                total = int(nch.sum().item())
                hub_of = torch.repeat_interleave(torch.arange(hubs.numel(), device=dev), nch)
                k = torch.arange(total, device=dev) - first[hub_of]
                row_start = self.rowptr[hubs].to(torch.int64)
                start = row_start[hub_of] + k * self.HUB_CHUNK
                end = torch.minimum(start + self.HUB_CHUNK, self.rowptr[hubs + 1].to(torch.int64)[hub_of])
                zc = torch.zeros_like(start)
                chunk_info = torch.stack([hub_of, start, end, zc], 1).to(torch.int32).contiguous()
                hub_info = torch.stack([hubs, first, nch, torch.zeros_like(nch)], 1).to(torch.int32).contiguous()
                tickets = torch.zeros(hubs.numel(), dtype=torch.int32, device=dev)
                self._plan = dict(n_hubs=int(hubs.numel()), n_chunks=total, hub_info=hub_info,
                                  chunk_info=chunk_info, tickets=tickets, partial=None)
        return self._plan or None

    # ---- SpMM ------------------------------------------------------------------------------------
    def spmm(self, x, mul=None, e1=None, e2=None, out=None, out2_mul=None, out2=None):
        """out = self · x  (gathered operand x ⊙ mul if mul is given; + e1 ⊙ e2 if given).
        out2_mul: also return out2 = out ⊙ out2_mul from the same launch (fp32, no `mul`); returns (out, out2)."""
        L = _ffi.lib()
        if x.dim() != 2 or x.shape[0] != self.n_cols:
            raise RuntimeError(f"spmm: x has shape {tuple(x.shape)}, adjacency is {self.n_rows}x{self.n_cols}")
        if x.dtype != self.val.dtype:
            raise RuntimeError(f"spmm: dtype mismatch ({x.dtype} features, {self.val.dtype} adjacency)")
        d = x.shape[1]
        if out is None:
            out = torch.empty(self.n_rows, d, dtype=x.dtype, device=x.device)
        rec = DeviceCSR.timer
        if rec is not None:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
        if out2_mul is not None:
            if x.dtype != torch.float32 or mul is not None:
                raise RuntimeError("spmm: the second output is fp32-only and excludes a gathered multiplier")
            if out2 is None:
                out2 = torch.empty_like(out)
            pl = self.plan()
            if pl is not None and (pl["partial"] is None or pl["partial"].shape[1] != d):
                pl["partial"] = torch.empty(pl["n_chunks"], d, dtype=torch.float32, device=x.device)
            _ffi.check(L.gdr_spmm_csr_dual_f32(
                _ffi.ptr(self.rowptr), _ffi.ptr(self.col), _ffi.ptr(self.val), self.n_rows, d, _ffi.ptr(x),
                _ffi.ptr(e1), _ffi.ptr(e2), _ffi.ptr(out), _ffi.ptr(out2_mul), _ffi.ptr(out2),
                self.HUB_THRESHOLD if pl else 0, pl["n_hubs"] if pl else 0, _ffi.ptr(pl["hub_info"]) if pl else None,
                pl["n_chunks"] if pl else 0, _ffi.ptr(pl["chunk_info"]) if pl else None,
                _ffi.ptr(pl["tickets"]) if pl else None, _ffi.ptr(pl["partial"]) if pl else None, _ffi.stream()),
                "spmm_csr_dual_f32")
        elif x.dtype == torch.float32:
            pl = self.plan()
            if pl is None:
                _ffi.check(L.gdr_spmm_csr_f32(_ffi.ptr(self.rowptr), _ffi.ptr(self.col), _ffi.ptr(self.val),
                                              self.n_rows, d, _ffi.ptr(x), _ffi.ptr(mul), _ffi.ptr(e1), _ffi.ptr(e2),
                                              _ffi.ptr(out), _ffi.stream()), "spmm_csr_f32")
            else:
                if pl["partial"] is None or pl["partial"].shape[1] != d:
                    pl["partial"] = torch.empty(pl["n_chunks"], d, dtype=torch.float32, device=x.device)
                _ffi.check(L.gdr_spmm_csr_planned_f32(
                    _ffi.ptr(self.rowptr), _ffi.ptr(self.col), _ffi.ptr(self.val), self.n_rows, d, _ffi.ptr(x),
                    _ffi.ptr(mul), _ffi.ptr(e1), _ffi.ptr(e2), _ffi.ptr(out), self.HUB_THRESHOLD, pl["n_hubs"],
                    _ffi.ptr(pl["hub_info"]), pl["n_chunks"], _ffi.ptr(pl["chunk_info"]), _ffi.ptr(pl["tickets"]),
                    _ffi.ptr(pl["partial"]), _ffi.stream()), "spmm_csr_planned_f32")
        elif x.dtype == torch.float64:
            if mul is not None or e1 is not None:
                raise RuntimeError("spmm: fused operands are fp32-only")
            _ffi.check(L.gdr_spmm_csr_f64(_ffi.ptr(self.rowptr), _ffi.ptr(self.col), _ffi.ptr(self.val), self.n_rows,
                                          d, _ffi.ptr(x), _ffi.ptr(out), _ffi.stream()), "spmm_csr_f64")
        else:
            raise RuntimeError(f"spmm: unsupported dtype {x.dtype}")
        if rec is not None:
            ev1.record()
            rec.append((ev0, ev1, self.nnz, self.n_rows, d, mul is not None, e1 is not None))
        if out2_mul is not None:
            return out, out2
        return out


def device_csr_of(adj):
    """DeviceCSR for the adjacency object the reference API passes around, built once and cached
    on the object (train.py passes the same tensor every iteration)."""
    if isinstance(adj, DeviceCSR):
        return adj
    cached = getattr(adj, "_gdr_csr", None)
    if cached is not None:
        return cached
    if not isinstance(adj, torch.Tensor):
        raise RuntimeError(f"adjacency must be a torch sparse tensor or DeviceCSR, got {type(adj)}")
    csr = DeviceCSR.from_torch_sparse(adj)
    try:
        adj._gdr_csr = csr
    except Exception:
        pass
    return csr
