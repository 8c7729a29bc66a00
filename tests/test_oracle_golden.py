"""CPU tests: the oracle restatement (oracle/*.py) against the golden fixtures that
oracle/gen_golden.py produced by running the reference itself.  No GPU, no /root/reference.

The fixtures were written by the same torch/numpy/scipy versions this image ships, so the oracle is
expected to reproduce them bit for bit; a tolerance (north_star's) is used only where a threaded
BLAS reduction order may legitimately differ between hosts.
"""
import io

import numpy as np
import pytest
import scipy.sparse as sp
import torch

import gdr_b200  # noqa: F401  (puts the hyphenated package on the path)
from gdr_b200 import synth

from .conftest import assert_close, load_golden
from oracle import gcn_oracle, rwr_oracle

GCN_CASES = ["gcn_knn257.npz", "gcn_knn600_l3.npz", "gcn_edge300_d256.npz"]


def _csr(g, prefix, shape):
    return sp.csr_matrix((g[prefix + "_data"], g[prefix + "_indices"], g[prefix + "_indptr"]), shape=shape)


@pytest.mark.parametrize("case", ["gcn_knn257.npz", "gcn_knn600_l3.npz"])
def test_gen_graph_oracle_matches_reference_golden(case):
    g = load_golden(case)
    n, d = g["embs"].shape
    X = g["embs"].astype(np.float64).T
    q_adj, qf, x_adj, xf = gcn_oracle.gen_graph(X[:, 0:3], X, 5, int(g["k"]))
    x_adj.sort_indices()
    assert np.array_equal(x_adj.indptr, g["x_adj_indptr"])
    assert np.array_equal(x_adj.indices, g["x_adj_indices"])
    np.testing.assert_allclose(x_adj.data, g["x_adj_data"], rtol=1e-14)
    q_adj.sort_indices()
    assert np.array_equal(q_adj.indices, g["q_adj_indices"])
    top, sim = gcn_oracle.topk_sets(X, int(g["k"]))
    assert np.array_equal(np.sort(top, axis=1), g["topk_sorted"])
    assert xf.shape == (n, d) and qf.shape == (3, d)
    # the adjacency is exactly symmetric with an empty diagonal
    assert (x_adj != x_adj.T).nnz == 0 and x_adj.diagonal().sum() == 0


@pytest.mark.parametrize("case", GCN_CASES)
def test_preprocess_graph_oracle_matches_reference_golden(case):
    g = load_golden(case)
    n = len(g["x_adj_indptr"]) - 1
    a_hat = gcn_oracle.preprocess_graph(_csr(g, "x_adj", (n, n)))
    a_hat.sort_indices()
    assert np.array_equal(a_hat.indptr, g["ahat_indptr"])
    assert np.array_equal(a_hat.indices, g["ahat_indices"])
    assert np.array_equal(a_hat.data, g["ahat_data"])           # fp64 bit-exact
    t = gcn_oracle.to_sparse_tensor(a_hat)
    assert np.array_equal(t._indices().numpy(), g["coo_indices"])
    assert np.array_equal(t._values().numpy(), g["coo_values"])


@pytest.mark.parametrize("case", GCN_CASES)
def test_training_oracle_matches_reference_golden(case):
    g = load_golden(case)
    n, d = g["embs"].shape
    adj = torch.sparse_coo_tensor(torch.from_numpy(g["coo_indices"]), torch.from_numpy(g["coo_values"]), (n, n))
    tr = gcn_oracle.Trainer(g["embs"], adj, d, num_layers=int(g["num_layers"]), layer_decay=float(g["layer_decay"]),
                            alpha=float(g["alpha"]), lr=float(g["lr"]), w_init=g["w_init"])
    sizes = g["batch_sizes"]
    offs = np.concatenate([[0], np.cumsum(sizes)])
    beta = float(g["beta"])
    losses = []
    for it in range(len(sizes)):
        b = torch.from_numpy(g["batches_flat"][offs[it]:offs[it + 1]])
        if it == 0:
            h0 = tr.forward().detach()
            assert_close(gcn_oracle.beta_percentile(h0, float(g["beta_percentile"])), g["beta"], what="beta")
        loss, h, grads = tr.step(b, beta)
        losses.append(loss.item())
        if it == 0:
            for nm, gr in zip(("dw1", "db1", "dw2", "db2"), grads):
                assert_close(gr.numpy(), g[f"{nm}_step0"], what=f"{nm} step0")
    assert_close(h.numpy(), g["h_last"], what="hidden_emb last")
    assert_close(np.array(losses), g["losses"], what="losses")
    for nm, p in zip(("w1", "b1", "w2", "b2"), tr.params):
        assert_close(p.detach().numpy(), g[f"{nm}_final"], what=f"{nm} final")
    assert_close(gcn_oracle.gss_loss(tr.forward().detach(), beta).item(), g["loss_all"], what="loss_all")
    buf = io.StringIO()
    np.savetxt(buf, g["h_last"][:2])
    assert buf.getvalue() == str(g["savetxt_head"])            # '%.18e', space separated, no header


def test_layer_oracle_intermediates():
    g = load_golden("gcn_knn257.npz")
    n, d = g["embs"].shape
    adj = torch.sparse_coo_tensor(torch.from_numpy(g["coo_indices"]), torch.from_numpy(g["coo_values"]), (n, n))
    w = torch.from_numpy(g["w_init"])
    b = torch.zeros(d)
    z, y, p, t = gcn_oracle.layer_forward(torch.from_numpy(g["embs"]), adj, w, b, w, b)
    assert_close(p.numpy(), g["layer1_P"], what="P")
    assert_close(t.numpy(), g["layer1_T"], what="T")
    assert_close(z.numpy(), g["layer1_Z"], what="Z")
    assert_close(y.numpy(), g["layer1_Y"], what="Y")


def test_rwr_oracle_matches_reference_golden():
    g = load_golden("rwr_msi600.npz")
    graph = synth.msi_synth(seed=int(g["seed"]), scale=float(g["scale"]))
    m0 = rwr_oracle.initial_matrix(graph["n"], graph["src"], graph["dst"], graph["weight"])
    sinks = [int(s) for s in graph["seeds"]]
    for row, ci in enumerate(g["chosen"]):
        s = int(graph["seeds"][ci])
        x, it = rwr_oracle.diffusion_profile(m0, s, graph["seed_proteins"], sinks, synth.RWR_ALPHA, synth.RWR_TOL,
                                             synth.RWR_MAX_ITER)
        assert it == int(g["iters"][row])
        np.testing.assert_allclose(x, g["profiles"][row], rtol=1e-13, atol=0)
        assert abs(x.sum() - 1.0) < 1e-9
        assert rwr_oracle.clean_file_name(graph["names"][s]) == str(g["file_names"][row])


def test_synth_generators_are_deterministic():
    a = synth.msi_synth(seed=7, scale=0.01)
    b = synth.msi_synth(seed=7, scale=0.01)
    assert np.array_equal(a["src"], b["src"]) and np.array_equal(a["weight"], b["weight"])
    # every directed edge has its reverse; weights are per-source typed (so generally asymmetric)
    fwd = set(zip(a["src"].tolist(), a["dst"].tolist()))
    assert all((v, u) in fwd for u, v in fwd)
    e = synth.synth_embeddings(50, 8, seed=3)
    assert e.dtype == np.float32 and e.shape == (50, 8)
    r, c = synth.rmat_coo(8, 4, 1, "cpu")
    assert r.numel() == 4 << 8 and int(r.max()) < 256 and int(c.max()) < 256
