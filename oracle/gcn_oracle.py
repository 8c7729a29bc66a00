"""ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's GCN hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package (gcn-drug-repurposing_b200) never does.

Parity status: PINNED AGAINST THE REFERENCE ITSELF.  The reference has no tests or golden
vectors for this path, so oracle/gen_golden.py imports the reference's own modules
from /root/reference in the build container, runs them on seeded inputs and (a) asserts that
every function below reproduces the reference output bit-for-bit (same library calls in the
same order), (b) writes the outputs to tests/golden/*.npz.  tests/test_oracle_golden.py re-checks
(b) on every run, without /root/reference.

The arithmetic lives in third-party libraries the reference calls (torch >=1.4 -> 2.11 here,
numpy, scipy); this restatement makes the same calls so that it times and rounds
like the reference does.  Each function cites the reference lines it follows.
"""
import numpy as np
import scipy.sparse as sp
import torch
import torch.nn.functional as F


# ---- graph construction -------------------------------------------------------------------
def gen_graph(Q, X, kq=5, k=5):
    """helpers/helper.py:25-58 (descriptor mode).  Q: d×nq, X: d×N fp64.  Returns
    (q_adj csr, Q.T, x_adj csr, X.T)."""
    q_sim = np.matmul(Q.T, X)                                   # :27
    q_top = np.argpartition(q_sim, -kq, 1)[:, -kq:]             # :32
    q_adj = np.zeros(q_sim.shape)
    for i in range(q_adj.shape[0]):                             # :35-36
        q_adj[i, q_top[i]] = q_sim[i, q_top[i]]
    q_adj = sp.csr_matrix(q_adj)
    x_sim = np.matmul(X.T, X)                                   # :39
    x_top = np.argpartition(x_sim, -k, 1)[:, -k:]               # :44
    x_adj = np.zeros(x_sim.shape)
    for i in range(x_adj.shape[0]):                             # :48-51
        x_adj[i, x_top[i]] = x_sim[i, x_top[i]]
        x_adj[x_top[i], i] = x_sim[i, x_top[i]]
        x_adj[i, i] = 0
    return q_adj, Q.T, sp.csr_matrix(x_adj), X.T


def topk_sets(X, k):
    """Per-row top-k index sets and the Gram values, as helper.py:39,44 computes them."""
    x_sim = np.matmul(X.T, X)
    return np.argpartition(x_sim, -k, 1)[:, -k:], x_sim


def preprocess_graph(adj):
    """helpers/helper.py:82-89."""
    adj_ = adj + sp.eye(adj.shape[0])
    rowsum = np.array(adj_.sum(1))
    d = sp.diags(np.power(rowsum, -0.5).flatten())
    return sp.csr_matrix(adj_.dot(d).transpose().dot(d).transpose())


def to_sparse_tensor(m):
    """helpers/helper.py:92-96 (modern constructor; same indices, values, flags)."""
    m = m.tocoo()
    i = torch.from_numpy(np.vstack([m.row, m.col]).astype(np.int64))
    v = torch.from_numpy(m.data.astype(np.float32))
    return torch.sparse_coo_tensor(i, v, torch.Size(m.shape))


# ---- model ----------------------------------------------------------------------------------
def init_weight(hidden_units, init_weights=1e-5):
    """modules/model.py:139-140 (consumes numpy RNG)."""
    w = np.random.randn(hidden_units, hidden_units) * init_weights
    w[np.where(np.eye(hidden_units) != 0)] = 1
    return torch.Tensor(w)


def layer_forward(features, adj, w1, b1, w2, b2):
    """modules/model.py:152-175.  Returns (pre_nonlinearity, output) and the intermediates."""
    ax = torch.sparse.mm(adj, features)                         # :163
    z1 = F.linear(ax, w1, b1)                                   # :165
    ax_x = torch.mul(ax, features)                              # :168
    t = torch.sparse.mm(adj, ax_x)                              # :169
    z2 = F.linear(t, w2, b2)                                    # :170
    z = z1 + z2                                                 # :172
    return z, F.elu(z), ax, t                                   # :173


def network_forward(x, adj, w1, b1, w2, b2, num_layers=2, layer_decay=0.3):
    """modules/model.py:197-207 (the same layer object at every depth)."""
    residual = None
    for _ in range(num_layers):
        pre, x = layer_forward(x, adj, w1, b1, w2, b2)[:2]
        if residual is not None:
            x = residual + layer_decay * x
        residual = pre
    return F.normalize(x, dim=1)


def gss_loss(embs, beta, alpha=1.0, index=None):
    """modules/model.py:214-221."""
    if index is not None:
        embs = embs[index]
    g = torch.mm(embs, embs.transpose(0, 1))
    return torch.mean(-0.5 * alpha * (F.relu(g) - beta) ** 2)


def beta_percentile(h0, q):
    """train.py:165-167."""
    h0 = h0.detach().cpu()
    return np.percentile(np.dot(h0, h0.transpose(0, 1)).flatten(), q)


class Trainer:
    """train.py:111-184 restated on explicit tensors: Adam(lr, wd 0) over (W1,b1,W2,b2)."""

    def __init__(self, features, adj, hidden_units, num_layers=2, layer_decay=0.3, init_weights=1e-5,
                 alpha=1.0, lr=3e-4, w_init=None):
        self.x = torch.as_tensor(features, dtype=torch.float32)
        self.adj = adj
        w0 = init_weight(hidden_units, init_weights) if w_init is None else torch.as_tensor(w_init).clone()
        self.w1 = w0.clone().requires_grad_(True)
        self.b1 = torch.zeros(hidden_units, requires_grad=True)
        self.w2 = w0.clone().requires_grad_(True)
        self.b2 = torch.zeros(hidden_units, requires_grad=True)
        self.params = [self.w1, self.b1, self.w2, self.b2]      # parameter order of the nn.Module
        self.num_layers, self.layer_decay, self.alpha = num_layers, layer_decay, alpha
        self.opt = torch.optim.Adam(self.params, lr=lr, weight_decay=0)

    def forward(self):
        return network_forward(self.x, self.adj, self.w1, self.b1, self.w2, self.b2, self.num_layers,
                               self.layer_decay)

    def step(self, batch, beta):
        """train.py:158-184 for one minibatch.  Returns (loss, hidden_emb, grads)."""
        h = self.forward()
        loss = gss_loss(h, beta, self.alpha, index=batch)
        self.opt.zero_grad()
        loss.backward()
        grads = [p.grad.detach().clone() for p in self.params]
        self.opt.step()
        return loss.detach(), h.detach(), grads


def savetxt_embs(path, h):
    """train.py:193."""
    np.savetxt(path, h)
