"""TEST INFRASTRUCTURE ONLY (imported by tests/, never by the product package).

CPU restatement of the Kipf & Welling GCN the reference vendors under multiscale/openne/gcn/.  The utils half
(`preprocess_features`, `normalize_adj`, `preprocess_adj`, `chebyshev_polynomials`) is numpy/scipy in the reference and is
PINNED: oracle/gen_golden_kipf.py imports the reference's own utils.py and asserts these restatements reproduce it bit for
bit.  The TF1 half (layers.py, models.py, metrics.py) cannot be executed here — TensorFlow 1 is not installable — so its
restatement below (torch CPU, fp32, autograd) is checked by reading only: PARITY UNPINNED for GraphConvolution / GCN /
masked_softmax_cross_entropy / masked_accuracy.
"""
import numpy as np
import scipy.sparse as sp
import torch


# ---- utils.py ---------------------------------------------------------------------------------------------
def sparse_to_tuple(mx):                                   # utils.py:66-80
    if not sp.isspmatrix_coo(mx):
        mx = mx.tocoo()
    return np.vstack((mx.row, mx.col)).transpose(), mx.data, mx.shape


def preprocess_features(features):                         # utils.py:83-91
    rowsum = np.array(features.sum(1))
    with np.errstate(divide="ignore"):
        r_inv = np.power(rowsum, -1).flatten()
    r_inv[np.isinf(r_inv)] = 0.
    features = sp.diags(r_inv).dot(sp.coo_matrix(features))
    return sparse_to_tuple(features)


def normalize_adj(adj):                                    # utils.py:94-101
    adj = sp.coo_matrix(adj)
    rowsum = np.array(adj.sum(1))
    with np.errstate(divide="ignore"):
        d_inv_sqrt = np.power(rowsum, -0.5).flatten()
    d_inv_sqrt[np.isinf(d_inv_sqrt)] = 0.
    d = sp.diags(d_inv_sqrt)
    return adj.dot(d).transpose().dot(d).tocoo()


def preprocess_adj(adj):                                   # utils.py:104-107
    return sparse_to_tuple(normalize_adj(adj + sp.eye(adj.shape[0])))


def chebyshev_polynomials(adj, k):                         # utils.py:123-147
    from scipy.sparse.linalg import eigsh
    adj_normalized = normalize_adj(adj)
    laplacian = sp.eye(adj.shape[0]) - adj_normalized
    largest_eigval, _ = eigsh(laplacian, 1, which='LM')
    scaled_laplacian = (2. / largest_eigval[0]) * laplacian - sp.eye(adj.shape[0])
    t_k = [sp.eye(adj.shape[0]), scaled_laplacian]
    for _ in range(2, k + 1):
        s_lap = sp.csr_matrix(scaled_laplacian, copy=True)
        t_k.append(2 * s_lap.dot(t_k[-1]) - t_k[-2])
    return [sparse_to_tuple(t) for t in t_k]


# ---- layers.py / metrics.py / models.py (torch CPU restatement; unpinned) --------------------------------------
def to_torch_sparse(t):
    coords, values, shape = t
    return torch.sparse_coo_tensor(torch.from_numpy(np.asarray(coords).T.astype(np.int64)),
                                   torch.from_numpy(np.asarray(values).astype(np.float32)), tuple(shape)).coalesce()


def graph_convolution(x, supports, weights, bias=None, act=torch.relu, sparse_inputs=False, featureless=False):
    """layers.py:165-190 with dropout 0: act(sum_i support_i · (x · W_i) + b)."""
    out = None
    for a, w in zip(supports, weights):
        if featureless:
            pre = w
        elif sparse_inputs:
            pre = torch.sparse.mm(x, w)
        else:
            pre = torch.mm(x, w)
        s = torch.sparse.mm(a, pre)
        out = s if out is None else out + s
    if bias is not None:
        out = out + bias
    return act(out) if act is not None else out


def masked_softmax_cross_entropy(preds, labels, mask):     # metrics.py:4-10
    loss = -(labels * torch.log_softmax(preds, dim=1)).sum(1)
    mask = mask.float()
    mask = mask / mask.mean()
    return (loss * mask).mean()


def masked_accuracy(preds, labels, mask):                  # metrics.py:13-21
    correct = (preds.argmax(1) == labels.argmax(1)).float()
    mask = mask.float()
    mask = mask / mask.mean()
    return (correct * mask).mean()


def gcn_forward(features, supports, w0, w1, sparse_inputs=True):
    """models.py:166-181 with dropout 0."""
    h = graph_convolution(features, supports, w0, act=torch.relu, sparse_inputs=sparse_inputs)
    return graph_convolution(h, supports, w1, act=None)


def gcn_loss(outputs, labels, mask, w0, weight_decay):     # models.py:148-156
    loss = 0.
    for w in w0:
        loss = loss + weight_decay * 0.5 * (w * w).sum()
    return loss + masked_softmax_cross_entropy(outputs, labels, mask)
