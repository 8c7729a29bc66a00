"""Drop-in mirror of the Kipf & Welling GCN the reference vendors inside OpenNE — SURVEY.md 8(f) row 3:

    GraphConvolution          multiscale/openne/gcn/layers.py:133-190     act(sum_i A_i·(X·W_i) + b)
    GCN                       multiscale/openne/gcn/models.py:134-181     two layers, weight decay on the first, Adam 0.01
    masked_softmax_cross_entropy / masked_accuracy       multiscale/openne/gcn/metrics.py:4-21
    preprocess_features / normalize_adj / preprocess_adj / chebyshev_polynomials / sparse_to_tuple / sample_mask
                              multiscale/openne/gcn/utils.py:17-21,66-147

Same class names, constructor arguments and `placeholders` keys as the TF1 original ('support', 'features', 'labels',
'labels_mask', 'dropout', 'num_features_nonzero'); the session/feed_dict machinery is replaced by torch modules whose
arithmetic runs in libgdr.so: support·(X·W) and sparse X·W on the SpMM kernel (spmm.cu), dense X·W / dW / dX on the
tcgen05 3xTF32 GEMM (gemm_tc.cu), bias + activation, activation backward + bias gradient, dropout and the masked
softmax cross-entropy in kipf.cu.  Parity: the utils functions are pinned against the reference's own utils.py
(oracle/gen_golden_kipf.py); the TF1 graph itself cannot be run here (TensorFlow 1 is not installable), so layers /
model / metrics are checked against a torch-CPU restatement (oracle/kipf_oracle.py) — "parity unpinned" for those.
There is no CPU path: CPU tensors raise.
"""
import numpy as np
import scipy.sparse as sp
import torch
import torch.nn as nn

from . import _ffi
from .graph import DeviceCSR


# ------------------------------------------------------------------------------------------------------------
# utils.py
# ------------------------------------------------------------------------------------------------------------
def sample_mask(idx, l):
    """utils.py:17-21."""
    mask = np.zeros(l, dtype=bool)
    mask[np.asarray(list(idx), dtype=np.int64)] = True
    return mask


def sparse_to_tuple(sparse_mx):
    """utils.py:66-80: (coords [nnz,2], values, shape) of a scipy matrix (or of every matrix in a list, in place)."""
    def one(mx):
        mx = mx if sp.isspmatrix_coo(mx) else mx.tocoo()
        return np.vstack((mx.row, mx.col)).transpose(), mx.data, mx.shape
    if isinstance(sparse_mx, list):
        for i in range(len(sparse_mx)):
            sparse_mx[i] = one(sparse_mx[i])
        return sparse_mx
    return one(sparse_mx)


def preprocess_features(features):
    """utils.py:83-91: rows scaled to sum 1 (empty rows stay empty), tuple representation."""
    features = sp.coo_matrix(features)
    rowsum = np.asarray(features.sum(1)).flatten()
    with np.errstate(divide="ignore"):
        r_inv = np.power(rowsum, -1.0)
    r_inv[np.isinf(r_inv)] = 0.0
    return sparse_to_tuple(sp.diags(r_inv).dot(features))


def normalize_adj(adj):
    """utils.py:94-101: (A·D^-1/2)ᵀ·D^-1/2 with D = rowsum(A), isolated rows -> 0; COO result."""
    adj = sp.coo_matrix(adj)
    rowsum = np.asarray(adj.sum(1)).flatten()
    with np.errstate(divide="ignore"):
        d = np.power(rowsum, -0.5)
    d[np.isinf(d)] = 0.0
    dm = sp.diags(d)
    return adj.dot(dm).transpose().dot(dm).tocoo()


def preprocess_adj(adj):
    """utils.py:104-107: normalize_adj(A + I) as a tuple."""
    return sparse_to_tuple(normalize_adj(adj + sp.eye(adj.shape[0])))


def chebyshev_polynomials(adj, k):
    """utils.py:123-147: T_0..T_k of the rescaled Laplacian 2L/lambda_max - I, tuple representation."""
    from scipy.sparse.linalg import eigsh
    print("Calculating Chebyshev polynomials up to order {}...".format(k))
    n = adj.shape[0]
    laplacian = sp.eye(n) - normalize_adj(adj)
    largest, _ = eigsh(laplacian, 1, which="LM")
    scaled = (2.0 / largest[0]) * laplacian - sp.eye(n)
    t_k = [sp.eye(n), scaled]
    for _ in range(2, k + 1):
        s_lap = sp.csr_matrix(scaled, copy=True)
        t_k.append(2 * s_lap.dot(t_k[-1]) - t_k[-2])
    return sparse_to_tuple(t_k)


# ------------------------------------------------------------------------------------------------------------
# device plumbing
# ------------------------------------------------------------------------------------------------------------
def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("gcn-drug-repurposing_b200 needs a CUDA device: there is no CPU implementation")
    return torch.device("cuda", torch.cuda.current_device())


def as_device_csr(m):
    """Adjacency / sparse features in any form the reference passes around -> fp32 DeviceCSR (cached on tensors)."""
    if isinstance(m, DeviceCSR):
        return m if m.val.dtype == torch.float32 else m.astype(torch.float32)
    if isinstance(m, torch.Tensor):
        from .graph import device_csr_of
        return device_csr_of(m)
    if isinstance(m, tuple) and len(m) == 3:   # (coords, values, shape): sparse_to_tuple / tf.sparse_placeholder feed
        coords, values, shape = m
        coords = np.asarray(coords)
        m = sp.coo_matrix((np.asarray(values), (coords[:, 0], coords[:, 1])), shape=tuple(shape))
    if sp.issparse(m):
        c = m.tocsr().astype(np.float32)
        c.sum_duplicates()
        c.sort_indices()
        return DeviceCSR.from_scipy(c, _device(), dtype=torch.float32)
    raise RuntimeError(f"expected a sparse matrix (scipy, tuple, torch sparse or DeviceCSR), got {type(m)}")


def _pad_cols(t, mult=4):
    r = (-t.shape[1]) % mult
    return t if r == 0 else torch.nn.functional.pad(t, (0, r))


def _gemm_nt(a, b):
    """a [m,k] · b [n,k]ᵀ -> [m,n] on the tcgen05 3xTF32 GEMM; k, n multiples of 4 (callers pad)."""
    L = _ffi.lib()
    a, b = a.contiguous(), b.contiguous()
    m, k = a.shape
    n = b.shape[0]
    c = torch.empty(m, n, dtype=torch.float32, device=a.device)
    ws = _ffi.workspace(L.gdr_gemm_nt_tc_workspace_bytes(m, n, k), a.device)
    _ffi.check(L.gdr_gemm_nt_tc_f32(_ffi.ptr(a), k, _ffi.ptr(b), k, m, n, k, _ffi.ptr(c), n, _ffi.ptr(ws), ws.numel(),
                                    _ffi.stream()), "gemm_nt_tc")
    return c


def _t_pad(t):
    """tᵀ as a contiguous [cols, rows4] tensor, rows padded with zeros to a multiple of 4 (K of the next GEMM)."""
    return _pad_cols(t.t().contiguous())


class _GraphConvFunction(torch.autograd.Function):
    """y = act(sum_i A_i · pre_i + b), pre_i = X·W_i (dense X), Xsp·W_i (sparse X) or W_i (featureless)."""

    @staticmethod
    def forward(ctx, x, x_csr, supports, bias, relu, featureless, *weights):
        L = _ffi.lib()
        out_dim = weights[0].shape[1]
        out = None
        wps = [_pad_cols(w.detach().float().contiguous()) for w in weights]
        xp = None
        if x is not None:
            xp = _pad_cols(x.detach().float().contiguous())          # K of X·W padded to a multiple of 4
        for a, wp in zip(supports, wps):
            if featureless:
                pre = wp
            elif x_csr is not None:
                pre = x_csr.spmm(wp)                                    # sparse X · W   (layers.py:175-176, sparse=True)
            else:
                pre = _gemm_nt(xp, _t_pad(wp))                          # X [n, in4] · (Wᵀ [outp, in4])ᵀ
            out = a.spmm(pre.contiguous(), e1=out, out=out) if out is not None else a.spmm(pre.contiguous())
        n, outp = out.shape
        bp = None
        if bias is not None:
            bp = torch.nn.functional.pad(bias.detach().float(), (0, outp - out_dim)).contiguous()
        y = torch.empty_like(out)
        _ffi.check(L.gdr_bias_act_f32(_ffi.ptr(out), _ffi.ptr(bp), n, outp, 1 if relu else 0, _ffi.ptr(y), _ffi.stream()),
                   "bias_act")
        ctx.supports, ctx.x_csr, ctx.relu, ctx.featureless = supports, x_csr, relu, featureless
        ctx.out_dim, ctx.has_bias, ctx.x_shape = out_dim, bias is not None, None if x is None else tuple(x.shape)
        ctx.save_for_backward(y, xp if xp is not None else torch.empty(0), *wps)
        return y[:, :out_dim] if outp != out_dim else y

    @staticmethod
    def backward(ctx, dy):
        L = _ffi.lib()
        y, xp, *wps = ctx.saved_tensors
        n, outp = y.shape
        dyp = _pad_cols(dy.float().contiguous()) if dy.shape[1] != outp else dy.float().contiguous()
        dz = torch.empty_like(y)
        db = torch.empty(outp, dtype=torch.float32, device=y.device) if ctx.has_bias else None
        _ffi.check(L.gdr_act_bwd_colsum_f32(_ffi.ptr(dyp), _ffi.ptr(y), n, outp, 1 if ctx.relu else 0, _ffi.ptr(dz),
                                            _ffi.ptr(db), _ffi.stream()), "act_bwd_colsum")
        need_dx = ctx.needs_input_grad[0] and ctx.x_shape is not None
        dx = None
        dws = []
        xt = _t_pad(xp) if (xp.numel() and not ctx.featureless and ctx.x_csr is None) else None   # Xᵀ [in4, n4]
        for a, wp in zip(ctx.supports, wps):
            dpre = a.transpose().spmm(dz)                               # A_iᵀ · dZ
            if ctx.featureless:
                dw = dpre
            elif ctx.x_csr is not None:
                dw = ctx.x_csr.transpose().spmm(dpre)                   # Xspᵀ · dPre
            else:
                dw = _gemm_nt(xt, _t_pad(dpre))                         # Xᵀ · dPre: [in4, outp]
                if need_dx:
                    part = _gemm_nt(dpre, wp if wp.shape[0] == xp.shape[1] else
                                    torch.nn.functional.pad(wp, (0, 0, 0, xp.shape[1] - wp.shape[0])))  # dPre · Wᵀ
                    dx = part if dx is None else dx + part
            dws.append(dw[:wp.shape[0], :ctx.out_dim].contiguous())
        if dx is not None:
            dx = dx[:, :ctx.x_shape[1]].contiguous()
        return (dx, None, None, db[:ctx.out_dim] if db is not None else None, None, None, *dws)


class _DropoutFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, keep):
        L = _ffi.lib()
        x = x.contiguous()
        u = torch.rand_like(x)                                          # the generator is plumbing; the mask rule is TF's
        y, m = torch.empty_like(x), torch.empty_like(x)
        _ffi.check(L.gdr_dropout_apply_f32(_ffi.ptr(x), _ffi.ptr(u), x.numel(), float(keep), _ffi.ptr(y), _ffi.ptr(m),
                                           _ffi.stream()), "dropout_apply")
        ctx.save_for_backward(m)
        return y

    @staticmethod
    def backward(ctx, dy):
        (m,) = ctx.saved_tensors
        return dy * m, None


def sparse_dropout(x_csr, keep_prob):
    """layers.py:22-29 on a DeviceCSR: entries survive with probability keep_prob and are scaled by 1/keep_prob.
    Dropped entries stay in the structure with value 0 (tf.sparse_retain removes them: same product)."""
    L = _ffi.lib()
    u = torch.rand_like(x_csr.val)
    v = torch.empty_like(x_csr.val)
    _ffi.check(L.gdr_dropout_apply_f32(_ffi.ptr(x_csr.val), _ffi.ptr(u), x_csr.val.numel(), float(keep_prob), _ffi.ptr(v),
                                       None, _ffi.stream()), "dropout_apply")
    return DeviceCSR(x_csr.rowptr, x_csr.col, v, x_csr.n_rows, x_csr.n_cols)


def glorot(shape, generator=None, device=None):
    """inits.py:12-17: U(-r, r), r = sqrt(6 / (fan_in + fan_out))."""
    r = float(np.sqrt(6.0 / (shape[0] + shape[1])))
    return (torch.rand(*shape, generator=generator, device=device) * 2.0 - 1.0) * r


_RELUS = (torch.relu, torch.nn.functional.relu)


class GraphConvolution(nn.Module):
    """layers.py:133-190.  `placeholders['support']` is the list of (normalised) adjacencies, `placeholders['dropout']`
    the drop probability used when the `dropout` argument is truthy; `act`: torch.relu (fused), None / identity, or any
    callable (applied with torch after the fused identity epilogue)."""

    def __init__(self, input_dim, output_dim, placeholders, dropout=0., sparse_inputs=False, act=torch.relu, bias=False,
                 featureless=False, **kwargs):
        super().__init__()
        allowed = {"name", "logging", "generator"}
        for k in kwargs:
            assert k in allowed, "Invalid keyword argument: " + k
        self.name = kwargs.get("name") or "graphconvolution"
        self.placeholders = placeholders
        self.use_dropout = bool(dropout)
        self.act = act
        self.support = [as_device_csr(s) for s in placeholders["support"]]
        self.sparse_inputs = sparse_inputs
        self.featureless = featureless
        gen = kwargs.get("generator")
        dev = _device()
        for i in range(len(self.support)):
            w = glorot([input_dim, output_dim], generator=gen).to(dev)
            setattr(self, "weights_" + str(i), nn.Parameter(w))
        self.bias = nn.Parameter(torch.zeros(output_dim, device=dev)) if bias else None

    @property
    def vars(self):
        d = {"weights_" + str(i): getattr(self, "weights_" + str(i)) for i in range(len(self.support))}
        if self.bias is not None:
            d["bias"] = self.bias
        return d

    def forward(self, inputs):
        p = float(self.placeholders.get("dropout", 0.0)) if self.use_dropout else 0.0
        keep = 1.0 - p
        x, x_csr = None, None
        if self.sparse_inputs:
            x_csr = as_device_csr(inputs)
            if p > 0.0 and self.training:
                x_csr = sparse_dropout(x_csr, keep)
        elif not self.featureless or inputs is not None:
            x = inputs
            if not x.is_cuda:
                raise RuntimeError("GraphConvolution: inputs must be CUDA tensors (no CPU implementation)")
            if p > 0.0 and self.training:
                x = _DropoutFunction.apply(x, keep)
        weights = [getattr(self, "weights_" + str(i)) for i in range(len(self.support))]
        fused_relu = self.act in _RELUS
        y = _GraphConvFunction.apply(x, x_csr, self.support, self.bias, fused_relu, self.featureless, *weights)
        if fused_relu or self.act is None:
            return y
        return self.act(y)


class _MaskedCEFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, mask):
        L = _ffi.lib()
        lg = logits.float().contiguous()
        n, c = lg.shape
        loss = torch.empty((), dtype=torch.float32, device=lg.device)
        acc = torch.empty((), dtype=torch.float32, device=lg.device)
        dl = torch.empty_like(lg)
        ws = _ffi.workspace(L.gdr_masked_softmax_ce_workspace_bytes(n), lg.device)
        _ffi.check(L.gdr_masked_softmax_ce_f32(_ffi.ptr(lg), c, _ffi.ptr(labels), _ffi.ptr(mask), n, c, _ffi.ptr(loss),
                                               _ffi.ptr(acc), _ffi.ptr(dl), _ffi.ptr(ws), ws.numel(), _ffi.stream()),
                   "masked_softmax_ce")
        ctx.save_for_backward(dl)
        ctx.mark_non_differentiable(acc)
        return loss, acc

    @staticmethod
    def backward(ctx, gloss, gacc):
        (dl,) = ctx.saved_tensors
        return dl * gloss, None, None


def _labels_and_mask(labels, mask, device):
    lb = torch.as_tensor(np.asarray(labels) if not isinstance(labels, torch.Tensor) else labels)
    mk = torch.as_tensor(np.asarray(mask) if not isinstance(mask, torch.Tensor) else mask)
    return lb.to(device=device, dtype=torch.float32).contiguous(), mk.to(device=device, dtype=torch.float32).contiguous()


def masked_softmax_cross_entropy(preds, labels, mask):
    """metrics.py:4-10."""
    lb, mk = _labels_and_mask(labels, mask, preds.device)
    return _MaskedCEFunction.apply(preds, lb, mk)[0]


def masked_accuracy(preds, labels, mask):
    """metrics.py:13-21."""
    lb, mk = _labels_and_mask(labels, mask, preds.device)
    with torch.no_grad():
        return _MaskedCEFunction.apply(preds.detach(), lb, mk)[1]


class GCN(nn.Module):
    """models.py:134-181: GraphConvolution(input_dim -> hidden1, relu, sparse inputs) + GraphConvolution(hidden1 ->
    classes, identity); loss = weight_decay * l2_loss(first layer's variables) + masked softmax cross-entropy; Adam with
    learning rate 0.01 (`self.optimizer`; torch.optim semantics — TF1's Adam adds epsilon outside the bias-corrected
    root, a difference below fp32 resolution for gradients far above 1e-8)."""

    def __init__(self, placeholders, input_dim, hidden1, weight_decay, **kwargs):
        super().__init__()
        self.placeholders = placeholders
        self.hidden1, self.weight_decay, self.input_dim = hidden1, weight_decay, input_dim
        self.output_dim = int(np.asarray(placeholders["labels"]).shape[1]) if not isinstance(
            placeholders["labels"], torch.Tensor) else int(placeholders["labels"].shape[1])
        gen = kwargs.get("generator")
        sparse_in = not isinstance(placeholders["features"], torch.Tensor) or placeholders["features"].is_sparse
        self.layers = nn.ModuleList([
            GraphConvolution(input_dim, hidden1, placeholders, act=torch.relu, dropout=True, sparse_inputs=sparse_in,
                             generator=gen),
            GraphConvolution(hidden1, self.output_dim, placeholders, act=None, dropout=True, generator=gen)])
        self._features = as_device_csr(placeholders["features"]) if sparse_in else placeholders["features"]
        from . import optim
        self.optimizer = optim.Adam(self.parameters(), lr=0.01)
        self.outputs = None

    def forward(self):
        h = self._features
        for layer in self.layers:
            h = layer(h)
        self.outputs = h
        return h

    def loss(self, outputs=None):
        out = self.forward() if outputs is None else outputs
        reg = 0.0
        for var in self.layers[0].vars.values():
            reg = reg + self.weight_decay * 0.5 * (var * var).sum()      # tf.nn.l2_loss = sum(t^2) / 2
        return reg + masked_softmax_cross_entropy(out, self.placeholders["labels"], self.placeholders["labels_mask"])

    def accuracy(self, outputs=None):
        out = self.outputs if outputs is None else outputs
        return masked_accuracy(out, self.placeholders["labels"], self.placeholders["labels_mask"])

    def predict(self):
        return torch.softmax(self.forward(), dim=1)

    def train_step(self):
        """One `sess.run([opt_op, loss, accuracy])` of the reference's training loop (openne/gcn/gcnAPI.py)."""
        self.train()
        out = self.forward()
        loss = self.loss(out)
        self.optimizer.zero_grad()
        loss.backward()
        self.optimizer.step()
        return loss.detach(), self.accuracy(out.detach())
