"""Drop-in mirror of the live classes of the reference's modules/model.py, backed by libgdr.so.

    GSS_GNNLayer                         modules/model.py:134-175
    ResidualGraphConvolutionalNetwork    modules/model.py:178-207
    GSS_loss                             modules/model.py:210-221

Same constructor arguments, same parameter names and order (gcn_layer.dense.weight, .dense.bias,
.dense2.weight, .dense2.bias), same forward signatures and return values, same RNG consumption at
construction, so `state_dict`, `model.parameters()` and an optimizer built on them are
interchangeable with the reference.  The arithmetic is hand-written sm_100a CUDA reached through
the C ABI (include/gdr.h); autograd sees each forward as one custom Function whose backward
launches the kernels autograd would otherwise derive op by op.  CUDA tensors only: a CPU tensor
raises (there is no CPU implementation on the product path).
"""
import os

import numpy as np
import torch
import torch.nn as nn
from torch.nn.parameter import Parameter

from . import _ffi
from .graph import device_csr_of

# Which kernel family runs the dense contractions (X·Wᵀ, dZ·W, dZᵀ·P):
#   "tc"   tcgen05 tensor cores, 3xTF32 split with fp32-level error (csrc/gemm_tc.cu)
#   "simt" exact-fp32 FFMA tiles (csrc/gemm_simt.cuh)
# Both are CUDA; this is a kernel choice, not a fallback.  Override with GDR_GEMM=simt|tc.
GEMM_BACKEND = os.environ.get("GDR_GEMM", "tc")

# The first layer's two aggregations P¹ = Â·X and T¹ = Â·(P¹ ⊙ X) (modules/model.py:163,168-169) depend only on the
# input features and the adjacency, not on the weights (SURVEY §3.1 "hoistable").  Reusing them is OPT-IN
# (`model.hoist_layer1 = True`; trainer.GCNTrainer does it for the resident feature matrix it owns): the owner of the
# features is the only one who can promise that nobody rewrites them behind torch's back.  The cache belongs to the
# model object (dropped with it) and is keyed on the feature tensor object, its `_version`, its shape and the
# adjacency's CSR object.  `_version` sees torch's in-place ops but NOT writes through `.data`, through raw pointers
# (libgdr's `out=` arguments) or edits of the CSR values: after any of those call `model.invalidate_layer1_cache()`.
# GDR_HOIST=0 switches the reuse off everywhere, whatever the models ask for.
HOIST_LAYER1 = os.environ.get("GDR_HOIST", "1") != "0"


# --------------------------------------------------------------------------------------------
# raw-tensor building blocks (no autograd)
# --------------------------------------------------------------------------------------------
def _check_dense(t, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: this path has no CPU implementation")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")
    return t.contiguous()


def _check_layer_shapes(x, w1, b1, w2, b2, csr):
    """The reference's nn.Linear(h, h) raises when hidden_units != feature width (model.py:142,165); the C ABI takes
    bare pointers, so the same condition has to be checked here."""
    if x.dim() != 2:
        raise RuntimeError(f"features must be [n, d], got shape {tuple(x.shape)}")
    n, d = x.shape
    for nm, w in (("dense.weight", w1), ("dense2.weight", w2)):
        if tuple(w.shape) != (d, d):
            raise RuntimeError(f"{nm} has shape {tuple(w.shape)} but the features are {n}x{d}: hidden_units must "
                               f"equal the feature width (the layer is Linear(h, h) applied to adj @ features)")
    for nm, b in (("dense.bias", b1), ("dense2.bias", b2)):
        if b.numel() != d:
            raise RuntimeError(f"{nm} has {b.numel()} elements, expected {d}")
    if csr.n_rows != n or csr.n_cols != n:
        raise RuntimeError(f"adjacency is {csr.n_rows}x{csr.n_cols} but there are {n} feature rows")
    for nm, t in (("dense.weight", w1), ("dense.bias", b1), ("dense2.weight", w2), ("dense2.bias", b2)):
        if not t.is_cuda or t.dtype != torch.float32:
            raise RuntimeError(f"{nm} must be an fp32 CUDA tensor (call model.cuda(); there is no CPU path)")


class Layer1Cache:
    """P¹, T¹ of one (feature tensor, adjacency) pair.  Holds a strong reference to the feature tensor so that its
    storage cannot be recycled under the same address while the entry lives."""

    def __init__(self):
        self.entry = None
        self.hits = 0

    def clear(self):
        self.entry = None

    def lookup(self, csr, key, x):
        c = self.entry
        if (c is not None and c[0] is csr and c[1]._cdata == key._cdata and c[2] == key._version
                and c[3] == tuple(x.shape)):      # same CSR, same TensorImpl, no torch-visible write since
            self.hits += 1
            return c[4], c[5]
        return None

    def store(self, csr, key, x, p, t):
        self.entry = (csr, key, key._version, tuple(x.shape), p, t)


def _aggregate(csr, x, hoist_key=None, cache=None):
    """P = Â·x and T = Â·(P ⊙ x).  hoist_key + cache: the caller's feature tensor OBJECT and the owning model's
    Layer1Cache when these are the first layer's aggregations and the model has opted in."""
    use = hoist_key is not None and cache is not None and HOIST_LAYER1
    if use:
        hit = cache.lookup(csr, hoist_key, x)
        if hit is not None:
            return hit
    p, px = csr.spmm(x, out2_mul=x)      # model.py:163, and torch.mul(P, X) of :168 from the same launch
    t = csr.spmm(px)                     # model.py:169
    if use:
        cache.store(csr, hoist_key, x, p, t)
    return p, t


def layer_forward_raw(csr, x, w1, b1, w2, b2, residual, decay, hoist_key=None, cache=None):
    """One GSS_GNNLayer application (+ the residual update of the network loop).
    Returns P = Â·x, T = Â·(P⊙x), Z (pre-nonlinearity), Xnext (= ELU(Z) or residual + decay*ELU(Z))."""
    L = _ffi.lib()
    _check_layer_shapes(x, w1, b1, w2, b2, csr)
    n, d = x.shape
    p, t = _aggregate(csr, x, hoist_key, cache)
    z = torch.empty_like(x)
    xnext = torch.empty_like(x)
    if GEMM_BACKEND == "tc":
        ws = _ffi.workspace(L.gdr_gcn_layer_fwd_tc_workspace_bytes(n, d), x.device)
        _ffi.check(L.gdr_gcn_layer_fwd_tc_f32(_ffi.ptr(p), _ffi.ptr(t), _ffi.ptr(w1), _ffi.ptr(b1), _ffi.ptr(w2),
                                              _ffi.ptr(b2), _ffi.ptr(residual), float(decay), n, d, _ffi.ptr(z),
                                              _ffi.ptr(xnext), _ffi.ptr(ws), ws.numel(), _ffi.stream()),
                   "gcn_layer_fwd_tc")
        return p, t, z, xnext
    _ffi.check(L.gdr_gcn_layer_fwd_f32(_ffi.ptr(p), _ffi.ptr(t), _ffi.ptr(w1), _ffi.ptr(b1), _ffi.ptr(w2),
                                       _ffi.ptr(b2), _ffi.ptr(residual), float(decay), n, d, _ffi.ptr(z),
                                       _ffi.ptr(xnext), _ffi.stream()), "gcn_layer_fwd")
    return p, t, z, xnext


def layer_backward_raw(csr_t, gz, x, p, t, w1, w2, grads, accumulate, need_gx, ws):
    """Backward of one layer given gZ (gradient w.r.t. the pre-nonlinearity).
    grads = (dW1, db1, dW2, db2) flat fp32 tensors that are written (accumulate=False) or added to."""
    L = _ffi.lib()
    n, d = x.shape
    dw1, db1, dw2, db2 = grads
    acc = 1 if accumulate else 0
    if GEMM_BACKEND == "tc":
        wsw = _ffi.workspace(L.gdr_linear_bwd_weight2_tc_workspace_bytes(n, d), x.device)
        _ffi.check(L.gdr_linear_bwd_weight2_tc_f32(_ffi.ptr(gz), _ffi.ptr(p), _ffi.ptr(t), n, d, acc, _ffi.ptr(dw1),
                                                   _ffi.ptr(db1), _ffi.ptr(dw2), _ffi.ptr(db2), _ffi.ptr(wsw),
                                                   wsw.numel(), _ffi.stream()), "linear_bwd_weight2_tc")
        if not need_gx:
            return None
        wsi = _ffi.workspace(L.gdr_linear_bwd_input_tc_workspace_bytes(n, d), x.device)
        gt = torch.empty_like(x)
        _ffi.check(L.gdr_linear_bwd_input_tc_f32(_ffi.ptr(gz), _ffi.ptr(w2), None, None, n, d, _ffi.ptr(gt),
                                                 _ffi.ptr(wsi), wsi.numel(), _ffi.stream()), "linear_bwd_input_tc(W2)")
        gr = csr_t.spmm(gt)
        gp = gt
        _ffi.check(L.gdr_linear_bwd_input_tc_f32(_ffi.ptr(gz), _ffi.ptr(w1), _ffi.ptr(gr), _ffi.ptr(x), n, d,
                                                 _ffi.ptr(gp), _ffi.ptr(wsi), wsi.numel(), _ffi.stream()),
                   "linear_bwd_input_tc(W1)")
        return csr_t.spmm(gp, e1=gr, e2=p)
    _ffi.check(L.gdr_linear_bwd_weight_f32(_ffi.ptr(gz), _ffi.ptr(p), n, d, acc, _ffi.ptr(dw1), _ffi.ptr(db1),
                                           _ffi.ptr(ws), ws.numel(), _ffi.stream()), "linear_bwd_weight(W1)")
    _ffi.check(L.gdr_linear_bwd_weight_f32(_ffi.ptr(gz), _ffi.ptr(t), n, d, acc, _ffi.ptr(dw2), _ffi.ptr(db2),
                                           _ffi.ptr(ws), ws.numel(), _ffi.stream()), "linear_bwd_weight(W2)")
    if not need_gx:
        return None
    gt = torch.empty_like(x)
    _ffi.check(L.gdr_linear_bwd_input_f32(_ffi.ptr(gz), _ffi.ptr(w2), None, None, n, d, _ffi.ptr(gt),
                                          _ffi.stream()), "linear_bwd_input(W2)")
    gr = csr_t.spmm(gt)                                   # dR = Âᵀ·dT
    gp = gt                                               # reuse the buffer: dP = dZ·W1 + dR ⊙ X
    _ffi.check(L.gdr_linear_bwd_input_f32(_ffi.ptr(gz), _ffi.ptr(w1), _ffi.ptr(gr), _ffi.ptr(x), n, d,
                                          _ffi.ptr(gp), _ffi.stream()), "linear_bwd_input(W1)")
    gx = csr_t.spmm(gp, e1=gr, e2=p)                      # dX = Âᵀ·dP + dR ⊙ P
    return gx


def _bwd_weight_workspace(n, d, device):
    return _ffi.workspace(_ffi.lib().gdr_linear_bwd_weight_workspace_bytes(n, d), device)


# --------------------------------------------------------------------------------------------
# autograd Functions
# --------------------------------------------------------------------------------------------
class _GSSLayerFunction(torch.autograd.Function):
    """(Z, Y) = layer(features) with Y = ELU(Z)."""

    @staticmethod
    def forward(ctx, features, w1, b1, w2, b2, csr):
        x = _check_dense(features, "features")
        p, t, z, y = layer_forward_raw(csr, x, w1.contiguous(), b1.contiguous(), w2.contiguous(), b2.contiguous(),
                                       None, 0.0)
        ctx.csr = csr
        ctx.save_for_backward(x, p, t, z, w1, w2)
        ctx.need_gx = features.requires_grad
        return z, y

    @staticmethod
    def backward(ctx, gz_direct, gy):
        x, p, t, z, w1, w2 = ctx.saved_tensors
        L = _ffi.lib()
        n, d = x.shape
        if gy is None:
            gy = torch.zeros_like(z)
        gz = torch.empty_like(z)
        gres = None if gz_direct is None else gz_direct.contiguous()
        _ffi.check(L.gdr_elu_bwd_f32(_ffi.ptr(gy.contiguous()), _ffi.ptr(z), _ffi.ptr(gres), 1.0, z.numel(),
                                     _ffi.ptr(gz), _ffi.stream()), "elu_bwd")
        dw1 = torch.empty_like(w1)
        dw2 = torch.empty_like(w2)
        db1 = torch.empty(d, dtype=torch.float32, device=x.device)
        db2 = torch.empty(d, dtype=torch.float32, device=x.device)
        ws = _bwd_weight_workspace(n, d, x.device)
        gx = layer_backward_raw(ctx.csr.transpose(), gz, x, p, t, w1.contiguous(), w2.contiguous(),
                                (dw1, db1, dw2, db2), False, ctx.need_gx, ws)
        return gx, dw1, db1, dw2, db2, None


class _ResGCNFunction(torch.autograd.Function):
    """hidden_emb = normalize(X^{L+1}) for the whole residual network (weights shared across depth)."""

    @staticmethod
    def forward(ctx, x_in, w1, b1, w2, b2, csr, num_layers, decay, cache=None):
        L = _ffi.lib()
        x = _check_dense(x_in, "x")
        w1c, b1c, w2c, b2c = w1.contiguous(), b1.contiguous(), w2.contiguous(), b2.contiguous()
        n, d = x.shape
        xs, ps, ts, zs = [], [], [], []
        residual = None
        cur = x
        for li in range(num_layers):
            # layer 1 aggregates the caller's features: reusable (when the model opted in: cache is not None) while
            # that tensor object is unchanged
            key = x_in if (li == 0 and cache is not None and x is x_in and not x_in.requires_grad) else None
            p, t, z, xnext = layer_forward_raw(csr, cur, w1c, b1c, w2c, b2c, residual, decay, hoist_key=key,
                                               cache=cache)
            xs.append(cur), ps.append(p), ts.append(t), zs.append(z)
            residual = z
            cur = xnext
        h = torch.empty_like(cur)
        norm = torch.empty(n, dtype=torch.float32, device=x.device)
        _ffi.check(L.gdr_row_normalize_fwd_f32(_ffi.ptr(cur), n, d, _ffi.ptr(h), _ffi.ptr(norm), _ffi.stream()),
                   "row_normalize_fwd")
        ctx.csr = csr
        ctx.num_layers = num_layers
        ctx.decay = float(decay)
        ctx.need_gx = x_in.requires_grad
        ctx.save_for_backward(h, norm, w1, w2, *xs, *ps, *ts, *zs)
        return h

    @staticmethod
    def backward(ctx, gh):
        L = _ffi.lib()
        nl = ctx.num_layers
        saved = ctx.saved_tensors
        h, norm, w1, w2 = saved[:4]
        xs = saved[4:4 + nl]
        ps = saved[4 + nl:4 + 2 * nl]
        ts = saved[4 + 2 * nl:4 + 3 * nl]
        zs = saved[4 + 3 * nl:4 + 4 * nl]
        n, d = h.shape
        dev = h.device
        w1c, w2c = w1.contiguous(), w2.contiguous()
        csr_t = ctx.csr.transpose()
        gxnext = torch.empty_like(h)
        _ffi.check(L.gdr_row_normalize_bwd_f32(_ffi.ptr(gh.contiguous()), _ffi.ptr(h), _ffi.ptr(norm), n, d,
                                               _ffi.ptr(gxnext), _ffi.stream()), "row_normalize_bwd")
        dw1 = torch.empty_like(w1)
        dw2 = torch.empty_like(w2)
        db1 = torch.empty(d, dtype=torch.float32, device=dev)
        db2 = torch.empty(d, dtype=torch.float32, device=dev)
        ws = _bwd_weight_workspace(n, d, dev)
        gres_above = None   # gradient reaching Z^i through the residual of layer i+1
        gx = None
        for i in range(nl, 0, -1):
            scale = ctx.decay if i >= 2 else 1.0
            gz = torch.empty_like(h)
            _ffi.check(L.gdr_elu_bwd_f32(_ffi.ptr(gxnext), _ffi.ptr(zs[i - 1]), _ffi.ptr(gres_above), scale,
                                         gz.numel(), _ffi.ptr(gz), _ffi.stream()), "elu_bwd")
            gres_above = gxnext if i >= 2 else None      # X^{i+1} = Z^{i-1} + decay*Y^i
            need_gx = (i >= 2) or ctx.need_gx
            gx = layer_backward_raw(csr_t, gz, xs[i - 1], ps[i - 1], ts[i - 1], w1c, w2c, (dw1, db1, dw2, db2),
                                    accumulate=(i != nl), need_gx=need_gx, ws=ws)
            gxnext = gx
        return (gx if ctx.need_gx else None), dw1, db1, dw2, db2, None, None, None, None


class _GSSLossFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, embs, index, alpha, beta):
        L = _ffi.lib()
        e = _check_dense(embs, "embs")
        n, d = e.shape
        if index is not None:
            if not index.is_cuda:
                raise RuntimeError("index must be a CUDA tensor (train.py:157 moves the batch with .cuda())")
            index = index.to(torch.int64).contiguous()
            b = int(index.numel())
        else:
            b = n
        loss = torch.empty((), dtype=torch.float32, device=e.device)
        ge = torch.empty(b, d, dtype=torch.float32, device=e.device)
        if GEMM_BACKEND == "tc":
            ws = _ffi.workspace(L.gdr_gss_loss_tc_workspace_bytes(b, d), e.device)
            _ffi.check(L.gdr_gss_loss_fwd_bwd_tc_f32(_ffi.ptr(e), _ffi.ptr(index), b, n, d, float(alpha), float(beta),
                                                     _ffi.ptr(loss), _ffi.ptr(ge), _ffi.ptr(ws), ws.numel(),
                                                     _ffi.stream()), "gss_loss_fwd_bwd_tc")
        else:
            ws = _ffi.workspace(L.gdr_gss_loss_workspace_bytes(b, d), e.device)
            _ffi.check(L.gdr_gss_loss_fwd_bwd_f32(_ffi.ptr(e), _ffi.ptr(index), b, n, d, float(alpha), float(beta),
                                                  _ffi.ptr(loss), _ffi.ptr(ge), _ffi.ptr(ws), ws.numel(),
                                                  _ffi.stream()), "gss_loss_fwd_bwd")
        ctx.index = index
        ctx.shape = (n, d)
        ctx.save_for_backward(ge)
        return loss

    @staticmethod
    def backward(ctx, gloss):
        L = _ffi.lib()
        (ge,) = ctx.saved_tensors
        n, d = ctx.shape
        b = ge.shape[0]
        gh = torch.zeros(n, d, dtype=torch.float32, device=ge.device)
        scale = gloss.to(torch.float32).reshape(1).contiguous()
        _ffi.check(L.gdr_scatter_rows_f32(_ffi.ptr(ge), _ffi.ptr(ctx.index), b, n, d, _ffi.ptr(scale), _ffi.ptr(gh),
                                          _ffi.stream()), "scatter_rows")
        return gh, None, None, None


# --------------------------------------------------------------------------------------------
# the reference's classes
# --------------------------------------------------------------------------------------------
class GSS_GNNLayer(nn.Module):
    """GNN base layer (modules/model.py:134-175): two Linear(h,h) sharing one initial matrix
    I + init_weights*randn (numpy RNG), zero biases."""

    def __init__(self, hidden_units, init_weights=1e-5):
        super().__init__()
        w0 = np.random.randn(hidden_units, hidden_units) * init_weights
        np.fill_diagonal(w0, 1.0)
        for name in ("dense", "dense2"):
            lin = nn.Linear(hidden_units, hidden_units)   # consumes torch RNG exactly as the reference does
            lin.weight = Parameter(torch.tensor(w0, dtype=torch.float32))
            lin.bias = Parameter(torch.zeros(hidden_units, dtype=torch.float32))
            setattr(self, name, lin)

    def forward(self, features, adj):
        csr = device_csr_of(adj)
        pre_nonlinearity, output = _GSSLayerFunction.apply(features, self.dense.weight, self.dense.bias,
                                                           self.dense2.weight, self.dense2.bias, csr)
        return pre_nonlinearity, output


class ResidualGraphConvolutionalNetwork(nn.Module):
    """modules/model.py:178-207.  forward(x, adj) -> hidden_emb (row-normalised)."""

    def __init__(self, train_batch_size, val_batch_size, num_layers=2, hidden_units=2048, init_weights=1e-5,
                 layer_decay=0.4):
        super().__init__()
        self.train_batch_size = train_batch_size
        self.val_batch_size = val_batch_size
        self.num_layers = num_layers
        self.hidden_units = hidden_units
        self.init_weights = init_weights
        self.layer_decay = layer_decay
        self.gcn_layer = GSS_GNNLayer(hidden_units, init_weights)
        # not in the reference: opt-in reuse of the first layer's aggregations (see HOIST_LAYER1 above)
        self.hoist_layer1 = False
        self._layer1_cache = Layer1Cache()

    def invalidate_layer1_cache(self):
        """Forget the cached P¹/T¹.  Needed after writing the features through `.data` / a raw pointer, or after
        editing the adjacency values in place: torch's version counter does not see those."""
        self._layer1_cache.clear()

    def decoder(self, x):
        """Unused by train.py (model.py:190-195); kept for API completeness, composed from this
        package's kernels: normalise, then relu(H·Hᵀ) densely (small inputs only)."""
        hidden_emb = _normalize_rows(x)
        adj_preds = torch.relu(hidden_emb @ hidden_emb.t())
        return adj_preds, hidden_emb

    def forward(self, x, adj):
        if self.num_layers < 1:
            # the reference would call F.normalize on the raw input
            return _normalize_rows(x)
        csr = device_csr_of(adj)
        g = self.gcn_layer
        return _ResGCNFunction.apply(x, g.dense.weight, g.dense.bias, g.dense2.weight, g.dense2.bias, csr,
                                     int(self.num_layers), float(self.layer_decay),
                                     self._layer1_cache if self.hoist_layer1 else None)


class _NormalizeFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x_in):
        L = _ffi.lib()
        x = _check_dense(x_in, "x")
        n, d = x.shape
        h = torch.empty_like(x)
        norm = torch.empty(n, dtype=torch.float32, device=x.device)
        _ffi.check(L.gdr_row_normalize_fwd_f32(_ffi.ptr(x), n, d, _ffi.ptr(h), _ffi.ptr(norm), _ffi.stream()),
                   "row_normalize_fwd")
        ctx.save_for_backward(h, norm)
        return h

    @staticmethod
    def backward(ctx, gh):
        L = _ffi.lib()
        h, norm = ctx.saved_tensors
        n, d = h.shape
        gx = torch.empty_like(h)
        _ffi.check(L.gdr_row_normalize_bwd_f32(_ffi.ptr(gh.contiguous()), _ffi.ptr(h), _ffi.ptr(norm), n, d,
                                               _ffi.ptr(gx), _ffi.stream()), "row_normalize_bwd")
        return gx


def _normalize_rows(x):
    return _NormalizeFunction.apply(x)


class GSS_loss():
    """modules/model.py:210-221 (gss_loss only; tri_loss / bpr_rdp_loss2 are broken stubs upstream)."""

    def __init__(self, alpha):
        self.alpha = alpha

    def gss_loss(self, embs, beta, index=None):
        """If index is None the loss runs over all rows."""
        return _GSSLossFunction.apply(embs, index, float(self.alpha), float(beta))
