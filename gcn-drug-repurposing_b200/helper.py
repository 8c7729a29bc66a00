"""Drop-in mirror of the graph-construction functions of helpers/helper.py that train.py uses:

    gen_graph                                   helpers/helper.py:25-58   (train.py:93)
    combine_graph                               helpers/helper.py:73-79   (train.py:97)
    preprocess_graph                            helpers/helper.py:82-89   (train.py:99-100)
    convert_sparse_matrix_to_sparse_tensor      helpers/helper.py:92-96   (train.py:101-104)

Same arguments and return types (scipy CSR in / out, torch sparse COO out) so the reference
driver runs unchanged; the arithmetic (N×N Gram + row top-k, the (A+I) normalisation, the
CSR conversions) runs on the GPU through libgdr.so.  The N×N similarity matrix is never
materialised: the Gram kernel keeps a running top-k per row.
"""
import time

import numpy as np
import scipy.sparse as sp
import torch

from . import _ffi
from .graph import DeviceCSR


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("gcn-drug-repurposing_b200 needs a CUDA device: there is no CPU implementation")
    return torch.device("cuda", torch.cuda.current_device())


GRAM_TOPK_TC_MAX_D = 256   # operand A of the tensor-core kernel lives in tensor memory (256 + d <= 512 columns)
GRAM_TOPK_TC_MAX_K = 16
last_gram_topk_info = {}   # backend and number of rows that took the exact fp64 path in the last gram_topk call


def gram_topk(x_rows, k, backend=None):
    """x_rows: [n, d] fp64 CUDA tensor.  Returns (idx int32 [n,k], val fp64 [n,k]) of the k largest
    entries of every row of x_rows·x_rowsᵀ (helpers/helper.py:39,44), by descending value.

    backend "tc" (default when d <= 256 and k <= 16, GDR_GEMM != simt): TF32 Gram tiles on tcgen05 with the
    top-k fused into the epilogue, exact fp64 re-rank of the candidates, certified per row, uncertified rows
    recomputed in fp64 (gdr_gram_topk_tc_f64).  backend "f64": the fp64 SIMT kernel (gdr_gram_topk_f64).
    Both give the same index sets and values, fp64 ties excepted."""
    L = _ffi.lib()
    n, d = x_rows.shape
    if backend is None:
        from . import model as _model
        backend = "tc" if (_model.GEMM_BACKEND == "tc" and d <= GRAM_TOPK_TC_MAX_D and k <= GRAM_TOPK_TC_MAX_K) else "f64"
    idx = torch.empty(n, k, dtype=torch.int32, device=x_rows.device)
    val = torch.empty(n, k, dtype=torch.float64, device=x_rows.device)
    last_gram_topk_info.clear()
    last_gram_topk_info["backend"] = backend
    if backend == "tc":
        ws = _ffi.workspace(L.gdr_gram_topk_tc_workspace_bytes(n, d, k), x_rows.device)
        nfb = torch.zeros(1, dtype=torch.int32, device=x_rows.device)
        _ffi.check(L.gdr_gram_topk_tc_f64(_ffi.ptr(x_rows), n, d, k, _ffi.ptr(idx), _ffi.ptr(val), _ffi.ptr(nfb),
                                          _ffi.ptr(ws), ws.numel(), _ffi.stream()), "gram_topk_tc_f64")
        last_gram_topk_info["fallback_rows"] = nfb   # device tensor: reading it synchronises, so leave that to the caller
    elif backend == "f64":
        _ffi.check(L.gdr_gram_topk_f64(_ffi.ptr(x_rows), n, d, k, _ffi.ptr(idx), _ffi.ptr(val), _ffi.stream()),
                   "gram_topk_f64")
    else:
        raise ValueError(f"gram_topk: unknown backend {backend!r}")
    return idx, val


def gram_percentile(h, q):
    """np.percentile(np.dot(h, h.T).flatten(), q) of train.py:165-167 for an fp32 CUDA tensor h [n, d],
    by radix select over recomputed Gram tiles (N×N is never stored).  Returns numpy.float32 like numpy."""
    import ctypes
    L = _ffi.lib()
    if not h.is_cuda or h.dtype != torch.float32:
        raise RuntimeError("gram_percentile: expected an fp32 CUDA tensor (no CPU implementation)")
    h = h.contiguous()
    n, d = h.shape
    out = ctypes.c_float(0.0)
    from . import model as _model
    if _model.GEMM_BACKEND == "tc":
        ws = _ffi.workspace(L.gdr_gram_percentile_tc_workspace_bytes(n, d), h.device)
        _ffi.check(L.gdr_gram_percentile_tc_f32(_ffi.ptr(h), n, d, float(q), ctypes.byref(out), _ffi.ptr(ws),
                                                ws.numel(), _ffi.stream()), "gram_percentile_tc_f32")
    else:
        ws = _ffi.workspace(L.gdr_gram_percentile_workspace_bytes(n, d), h.device)
        _ffi.check(L.gdr_gram_percentile_f32(_ffi.ptr(h), n, d, float(q), ctypes.byref(out), _ffi.ptr(ws), ws.numel(),
                                             _ffi.stream()), "gram_percentile_f32")
    return np.float32(out.value)


def knn_adjacency(x_rows, k):
    """Symmetrised kNN adjacency of helpers/helper.py:39-53 as a scipy CSR (fp64, empty diagonal):
    x_adj[i, j] = x_adj[j, i] = <x_i, x_j> for j in top-k(i); the diagonal is cleared."""
    n = x_rows.shape[0]
    idx, val = gram_topk(x_rows, k)
    rows = torch.arange(n, device=x_rows.device, dtype=torch.int64).repeat_interleave(k)
    cols = idx.reshape(-1).to(torch.int64)
    vals = val.reshape(-1)
    keep = rows != cols                                    # x_adj[i, i] = 0 (helper.py:51)
    rows, cols, vals = rows[keep], cols[keep], vals[keep]
    # both directions; an edge selected from both ends carries the same value (S is symmetric),
    # so duplicates are dropped rather than summed (the reference overwrites, helper.py:49-50)
    r2 = torch.cat([rows, cols])
    c2 = torch.cat([cols, rows])
    v2 = torch.cat([vals, vals])
    csr = DeviceCSR.from_coo(r2, c2, v2, n, n)
    rowptr = csr.rowptr.cpu().numpy().astype(np.int64)
    col = csr.col.cpu().numpy()
    v = csr.val.cpu().numpy()
    # drop duplicate (row, col) pairs (sorted, so duplicates are adjacent); explicit zeros vanish
    # in sp.csr_matrix(dense) upstream, so drop them here too
    row_of = np.repeat(np.arange(n), np.diff(rowptr))
    first = np.ones(len(col), dtype=bool)
    if len(col) > 1:
        first[1:] = (row_of[1:] != row_of[:-1]) | (col[1:] != col[:-1])
    first &= v != 0.0
    m = sp.csr_matrix((v[first], (row_of[first], col[first])), shape=(n, n))
    m.sort_indices()
    return m


def gen_graph(Q, X, kq=5, k=5, q_RANSAC_graph=None, x_RANSAC_graph=None):
    """Q: d×nq, X: d×N (columns are nodes, fp64 as train.py:80-82 produces).  Returns
    (q_adj, Q.T, x_adj, X.T) exactly as the reference: q_adj nq×N CSR of the top-kq similarities of
    every query, x_adj the symmetrised N×N kNN graph with an empty diagonal."""
    t0 = time.time()
    dev = _device()
    Q = np.asarray(Q, dtype=np.float64)
    X = np.asarray(X, dtype=np.float64)
    n = X.shape[1]
    x_rows = torch.from_numpy(np.ascontiguousarray(X.T)).to(dev)

    # --- query graph (3 rows in train.py; vestigial but part of the return value) -------------
    q_sim = (torch.from_numpy(np.ascontiguousarray(Q.T)).to(dev) @ x_rows.t())  # nq×N, tiny
    if q_RANSAC_graph is not None:
        q_top = torch.as_tensor(np.asarray(q_RANSAC_graph)[:, 0:kq], device=dev, dtype=torch.int64)
    else:
        q_top = torch.topk(q_sim, kq, dim=1).indices
    q_vals = torch.gather(q_sim, 1, q_top)
    nq = q_sim.shape[0]
    q_rows = torch.arange(nq, device=dev).repeat_interleave(q_top.shape[1])
    q_adj = sp.csr_matrix((q_vals.reshape(-1).cpu().numpy(),
                           (q_rows.cpu().numpy(), q_top.reshape(-1).cpu().numpy())), shape=(nq, n))
    q_adj.eliminate_zeros()
    q_adj.sort_indices()

    # --- index graph ------------------------------------------------------------------------------
    if x_RANSAC_graph is not None:
        # helpers/helper.py:41-42 takes the first k columns of a precomputed retrieval graph; those .npy files belong
        # to the image-retrieval datasets the reference was forked from and are absent upstream too
        raise NotImplementedError("ransac graph modes need the retrieval datasets (absent upstream too)")
    x_adj = knn_adjacency(x_rows, k)

    print('Created G_q with [kq={}] [shape=[{}, {}]] , G with [k={}] [shape=[{}, {}]] in {:.2f}s'.format(
        kq, q_adj.shape[0], q_adj.shape[1], k, x_adj.shape[0], x_adj.shape[1], time.time() - t0))
    return q_adj, Q.T, x_adj, X.T


def combine_graph(q_adj, x_adj):
    """helpers/helper.py:73-79: stack the query rows on top of the index graph and prepend nq
    empty columns (host-side scipy; three rows of glue, not on the hot path)."""
    stacked = sp.vstack((q_adj, x_adj))
    pad = sp.csr_matrix((stacked.shape[0], q_adj.shape[0]))
    return sp.csr_matrix(sp.hstack((pad, stacked)))


def preprocess_graph(adj):
    """D^-1/2 (A+I) D^-1/2 with D = rowsum(A+I), fp64, returned as scipy CSR with sorted indices."""
    dev = _device()
    a = sp.csr_matrix(adj)
    if a.dtype != np.float64:
        a = a.astype(np.float64)
    a.sum_duplicates()
    a.sort_indices()
    csr = DeviceCSR.from_scipy(a, dev, dtype=torch.float64)
    out = csr.normalized_with_identity()
    return out.to_scipy()


def convert_sparse_matrix_to_sparse_tensor(X):
    """scipy sparse -> torch sparse COO (int64 indices [2,nnz] row-major, fp32 values, on the CPU like the
    reference; train.py:120 moves it with .cuda())."""
    coo = X.tocoo()
    i = torch.from_numpy(np.vstack([coo.row, coo.col]).astype(np.int64))
    v = torch.from_numpy(coo.data.astype(np.float32))
    return torch.sparse_coo_tensor(i, v, torch.Size(coo.shape))
