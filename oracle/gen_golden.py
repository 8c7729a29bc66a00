"""Generate tests/golden/*.npz by running the REFERENCE ITSELF (imported from /root/reference, which
exists only in the build container) on seeded synthetic inputs, and pin the oracle restatement
(oracle/gcn_oracle.py, oracle/rwr_oracle.py) against it bit for bit.

    python oracle/gen_golden.py            # rewrites tests/golden/, exits non-zero on any mismatch

TEST INFRASTRUCTURE ONLY.  Nothing here is imported by the product package.
"""
import os
import sys
import warnings

import numpy as np
import scipy
import scipy.sparse as sp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.path.insert(0, ROOT)
warnings.filterwarnings("ignore")

from oracle import gcn_oracle, rwr_oracle  # noqa: E402
import importlib  # noqa: E402

synth = importlib.import_module("gcn-drug-repurposing_b200.synth")
GOLD = os.path.join(ROOT, "tests", "golden")


def import_reference():
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "multiscale"))
    import networkx as nx
    # the five shims of SURVEY §0.5 (APIs removed from networkx 3 / scipy >= 1.12)
    if not hasattr(nx, "to_scipy_sparse_matrix"):
        nx.to_scipy_sparse_matrix = lambda g, nodelist=None, weight="weight", dtype=None: sp.csr_matrix(
            nx.to_scipy_sparse_array(g, nodelist=nodelist, weight=weight, dtype=dtype))
    counters = {"absolute": 0}

    def counting_absolute(a):
        counters["absolute"] += 1
        return np.absolute(a)

    scipy.array, scipy.where, scipy.repeat = np.array, np.where, np.repeat
    scipy.absolute = counting_absolute
    import modules.model as ref_model
    import helpers.helper as ref_helper
    from diff_prof.diffusion_profiles import DiffusionProfiles
    return ref_model, ref_helper, DiffusionProfiles, counters


def same(a, b, what):
    a, b = np.asarray(a), np.asarray(b)
    if a.shape != b.shape or not np.array_equal(a, b, equal_nan=True):
        raise SystemExit(f"ORACLE != REFERENCE at {what}: max|diff|="
                         f"{np.max(np.abs(a.astype(np.float64) - b.astype(np.float64))) if a.shape == b.shape else 'shape'}")


def csr_parts(m):
    m = sp.csr_matrix(m)
    m.sort_indices()
    return m.indptr.astype(np.int32), m.indices.astype(np.int32), m.data.astype(np.float64)


def gcn_case(name, ref_model, ref_helper, n, d, k, batch, num_layers, seed, adjacency="knn", steps=3, lean=False):
    embs = synth.synth_embeddings(n, d, seed=seed)
    X = embs.astype(np.float64).T
    Q = X[:, 0:3]
    out = {"embs": embs, "k": k, "num_layers": num_layers, "layer_decay": 0.3, "alpha": 1.0, "lr": 3e-4,
           "beta_percentile": 98.0}
    if adjacency == "knn":
        q_adj, qf, x_adj, xf = ref_helper.gen_graph(Q, X, 5, k)
        o_q_adj, _, o_x_adj, _ = gcn_oracle.gen_graph(Q, X, 5, k)
        for a, b, w in zip(csr_parts(x_adj), csr_parts(o_x_adj), ("indptr", "indices", "data")):
            same(a, b, f"{name}.gen_graph.x_adj.{w}")
        for a, b, w in zip(csr_parts(q_adj), csr_parts(o_q_adj), ("indptr", "indices", "data")):
            same(a, b, f"{name}.gen_graph.q_adj.{w}")
        top, sim = gcn_oracle.topk_sets(X, k)
        out["topk_sorted"] = np.sort(top, axis=1).astype(np.int32)
        # gap between the k-th and (k+1)-th similarity of every row: rows with a tiny gap are "ties"
        part = np.sort(sim, axis=1)
        out["topk_gap"] = (part[:, -k] - part[:, -k - 1])
        out["q_adj_indptr"], out["q_adj_indices"], out["q_adj_data"] = csr_parts(q_adj)
    else:
        g = synth.msi_synth(seed=seed, scale=n / 29959.0)
        assert g["n"] >= n
        keep = (g["src"] < n) & (g["dst"] < n)
        x_adj = sp.csr_matrix((g["weight"][keep], (g["src"][keep], g["dst"][keep])), shape=(n, n))
    out["x_adj_indptr"], out["x_adj_indices"], out["x_adj_data"] = csr_parts(x_adj)
    a_hat = ref_helper.preprocess_graph(x_adj)
    same(csr_parts(a_hat)[2], csr_parts(gcn_oracle.preprocess_graph(x_adj))[2], f"{name}.preprocess_graph")
    out["ahat_indptr"], out["ahat_indices"], out["ahat_data"] = csr_parts(a_hat)
    adj_t = ref_helper.convert_sparse_matrix_to_sparse_tensor(a_hat)
    o_adj_t = gcn_oracle.to_sparse_tensor(a_hat)
    same(adj_t._indices().numpy(), o_adj_t._indices().numpy(), f"{name}.coo.indices")
    same(adj_t._values().numpy(), o_adj_t._values().numpy(), f"{name}.coo.values")
    out["coo_indices"] = adj_t._indices().numpy()
    out["coo_values"] = adj_t._values().numpy()

    # --- model: reference classes vs oracle functions, identical RNG streams ----------------------
    torch.manual_seed(seed)
    np.random.seed(seed)
    model = ref_model.ResidualGraphConvolutionalNetwork(batch, n, num_layers=num_layers, hidden_units=d,
                                                        init_weights=1e-5, layer_decay=0.3)
    np.random.seed(seed)
    trainer = gcn_oracle.Trainer(embs, o_adj_t, d, num_layers=num_layers, layer_decay=0.3, lr=3e-4)
    same(model.gcn_layer.dense.weight.detach().numpy(), trainer.w1.detach().numpy(), f"{name}.init_w")
    out["w_init"] = model.gcn_layer.dense.weight.detach().numpy().copy()
    names = [n_ for n_, _ in model.named_parameters()]
    assert names == ["gcn_layer.dense.weight", "gcn_layer.dense.bias", "gcn_layer.dense2.weight",
                     "gcn_layer.dense2.bias"], names
    feats = torch.tensor(embs, dtype=torch.float32)
    opt = torch.optim.Adam(model.parameters(), lr=3e-4, weight_decay=0)
    loss_fcn = ref_model.GSS_loss(1.0).gss_loss

    # layer-level intermediates of the first layer (reference GSS_GNNLayer.forward)
    pre, y = model.gcn_layer(feats, adj_t)
    o_pre, o_y, o_p, o_t = gcn_oracle.layer_forward(feats, o_adj_t, trainer.w1, trainer.b1, trainer.w2, trainer.b2)
    same(pre.detach().numpy(), o_pre.detach().numpy(), f"{name}.layer.pre")
    same(y.detach().numpy(), o_y.detach().numpy(), f"{name}.layer.out")
    if not lean:
        out["layer1_P"], out["layer1_T"] = o_p.detach().numpy(), o_t.detach().numpy()
        out["layer1_Z"], out["layer1_Y"] = pre.detach().numpy(), y.detach().numpy()

    g = torch.Generator().manual_seed(seed)
    perm = torch.randperm(n, generator=g)
    batches = [perm[i * batch:(i + 1) * batch] for i in range(steps)]
    out["batches_flat"] = np.concatenate([b.numpy() for b in batches])
    out["batch_sizes"] = np.array([len(b) for b in batches], dtype=np.int64)
    beta = None
    losses, hs, grads = [], [], []
    for it, b in enumerate(batches):
        h = model(x=feats, adj=adj_t)
        if it == 0:
            h0 = h.cpu().data
            beta = np.percentile(np.dot(h0, h0.transpose(0, 1)).flatten(), 98.0)   # train.py:166-167
            same(beta, gcn_oracle.beta_percentile(h0, 98.0), f"{name}.beta")
        loss = loss_fcn(embs=h, beta=beta, index=b.detach())
        opt.zero_grad()
        loss.backward()
        ref_grads = [p.grad.detach().numpy().copy() for p in model.parameters()]
        opt.step()
        o_loss, o_h, o_grads = trainer.step(b, beta)
        same(h.detach().numpy(), o_h.numpy(), f"{name}.step{it}.hidden_emb")
        same(loss.detach().numpy(), o_loss.numpy(), f"{name}.step{it}.loss")
        for a_, b_, w in zip(ref_grads, o_grads, names):
            same(a_, b_.numpy(), f"{name}.step{it}.grad.{w}")
        losses.append(float(loss))
        hs.append(h.detach().numpy().copy())
        grads.append(ref_grads)
    for p_ref, p_or, w in zip(model.parameters(), trainer.params, names):
        same(p_ref.detach().numpy(), p_or.detach().numpy(), f"{name}.final.{w}")
    out["beta"] = np.float32(beta)
    out["losses"] = np.array(losses, dtype=np.float64)
    if not lean:
        out["h_step0"] = hs[0]
    out["h_last"] = hs[-1]
    for w, g0, g2 in zip(("dw1", "db1", "dw2", "db2"), grads[0], grads[-1]):
        out[f"{w}_step0"] = g0
        out[f"{w}_last"] = g2
    for w, p in zip(("w1", "b1", "w2", "b2"), model.parameters()):
        out[f"{w}_final"] = p.detach().numpy().copy()
    # loss on all rows (index=None)
    h_all = model(x=feats, adj=adj_t)
    out["loss_all"] = np.float64(loss_fcn(embs=h_all, beta=beta).item())
    # the output text format (train.py:193): first row as text
    import io
    buf = io.StringIO()
    np.savetxt(buf, hs[-1][:2])
    out["savetxt_head"] = np.array(buf.getvalue())
    np.savez_compressed(os.path.join(GOLD, f"gcn_{name}.npz"), **out)
    print(f"gcn_{name}: n={n} d={d} nnz(Â)={a_hat.nnz} beta={beta:.6f} losses={losses}")


class FakeMSI:
    """The attributes DiffusionProfiles reads from an MSI object (msi.py:148-205)."""

    def __init__(self, g):
        import networkx as nx
        names = g["names"]
        self.graph = nx.DiGraph()
        self.graph.add_nodes_from(names)
        for u, v, w in zip(g["src"], g["dst"], g["weight"]):
            self.graph.add_edge(names[u], names[v], weight=float(w))
        self.nodelist = list(names)
        self.node2idx = {nm: i for i, nm in enumerate(names)}
        t = g["node_type"]
        self.drugs_in_graph = [names[i] for i in g["seeds"] if t[i] == synth.DRUG]
        self.indications_in_graph = [names[i] for i in g["seeds"] if t[i] == synth.INDICATION]
        self.drug_or_indication2proteins = {
            names[s]: set(names[p] for p in g["seed_proteins"].get(int(s), [])) for s in g["seeds"]}


def rwr_case(name, DiffusionProfiles, counters, scale, seed, n_seeds):
    g = synth.msi_synth(seed=seed, scale=scale)
    msi = FakeMSI(g)
    dp = DiffusionProfiles(alpha=synth.RWR_ALPHA, max_iter=synth.RWR_MAX_ITER, tol=synth.RWR_TOL,
                           weights=None, num_cores=1, save_load_file_path=None)
    dp.get_initial_M(msi)
    m0 = rwr_oracle.initial_matrix(g["n"], g["src"], g["dst"], g["weight"])
    same(dp.initial_M.toarray(), m0.toarray(), f"{name}.initial_M")
    rng = np.random.default_rng(seed)
    chosen = np.sort(rng.choice(len(g["seeds"]), size=min(n_seeds, len(g["seeds"])), replace=False))
    profiles, iters = [], []
    sinks = [int(s) for s in g["seeds"]]
    for ci in chosen:
        s = int(g["seeds"][ci])
        nm = g["names"][s]
        M = dp.convert_M_to_make_all_drugs_indications_sinks_except_selected(msi, [nm])
        M, S = dp.refine_M_S(M)
        per = dp.get_personalization_dictionary([nm], msi.nodelist)
        counters["absolute"] = 0
        x = dp.power_iteration(M, S, msi.nodelist, per)
        it = counters["absolute"]
        ox, oit = rwr_oracle.diffusion_profile(m0, s, g["seed_proteins"], sinks, synth.RWR_ALPHA, synth.RWR_TOL,
                                               synth.RWR_MAX_ITER)
        same(x, ox, f"{name}.profile[{nm}]")
        assert it == oit, (nm, it, oit)
        profiles.append(x), iters.append(it)
    np.savez_compressed(os.path.join(GOLD, f"rwr_{name}.npz"), scale=scale, seed=seed, chosen=chosen,
                        profiles=np.stack(profiles), iters=np.array(iters, dtype=np.int32),
                        file_names=np.array([dp.clean_file_name(g["names"][int(g["seeds"][c])]) for c in chosen]))
    print(f"rwr_{name}: n={g['n']} nnz={len(g['src'])} seeds={len(chosen)} iters {min(iters)}..{max(iters)}")


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref_model, ref_helper, DiffusionProfiles, counters = import_reference()
    gcn_case("knn257", ref_model, ref_helper, n=257, d=128, k=5, batch=64, num_layers=2, seed=1)
    gcn_case("knn600_l3", ref_model, ref_helper, n=600, d=128, k=5, batch=200, num_layers=3, seed=2, lean=True)
    gcn_case("edge300_d256", ref_model, ref_helper, n=300, d=256, k=5, batch=128, num_layers=2, seed=3,
             adjacency="edge", lean=True)
    rwr_case("msi600", DiffusionProfiles, counters, scale=0.02, seed=5, n_seeds=48)
    gcn_case("knn2000", ref_model, ref_helper, n=2000, d=128, k=5, batch=512, num_layers=2, seed=4, lean=True)
    print("golden fixtures written; oracle == reference on every compared tensor")


if __name__ == "__main__":
    main()
