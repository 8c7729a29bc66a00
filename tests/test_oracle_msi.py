"""CPU tests: the oracle against the REAL-interactome fixtures (tests/golden/msi_*_graph.npz, rwr_msi_*.npz,
gcn_msi_noppi_edge.npz, rank_covid.npz, gcn_knn2000.npz) that oracle/gen_golden_msi.py / gen_golden.py produced by
running the reference itself.  No GPU, no /root/reference.  Sized to run in about a minute."""
import numpy as np
import pytest
import scipy.sparse as sp
import torch

import gdr_b200  # noqa: F401
from gdr_b200 import synth

from .conftest import assert_close, load_golden
from .msi_fixture import GoldenMSI
from oracle import gcn_oracle, rank_oracle, rwr_oracle


@pytest.mark.parametrize("name,n,nnz,n_seeds", [("msi_noppi", 22550, 182204, 2501), ("msi_covid", 20228, 132392, 1662)])
def test_real_msi_fixture_dimensions(name, n, nnz, n_seeds):
    g = load_golden(f"{name}_graph.npz")                       # SURVEY §0.4 [probed] figures
    assert len(g["names"]) == n and len(g["src"]) == nnz and len(g["seeds"]) == n_seeds
    fwd = set(zip(g["src"].tolist(), g["dst"].tolist()))
    assert all((v, u) in fwd for u, v in list(fwd)[:5000])     # every edge stored in both directions
    r = load_golden(f"rwr_{name}.npz")
    assert len(r["iters"]) >= 64 and r["profiles_full"].shape[1] == n


@pytest.mark.parametrize("name", ["msi_noppi", "msi_covid"])
def test_rwr_oracle_on_real_msi(name):
    g = load_golden(f"{name}_graph.npz")
    r = load_golden(f"rwr_{name}.npz")
    n = len(g["names"])
    m0 = rwr_oracle.initial_matrix(n, g["src"], g["dst"], g["weight"])
    sp_idx = {int(s): g["seed_protein_idx"][g["seed_protein_ptr"][i]:g["seed_protein_ptr"][i + 1]]
              for i, s in enumerate(g["seeds"])}
    sinks = [int(s) for s in g["seeds"]]
    for row in (0, 5, 11, 40, 63):                              # 5 of the 64 seeds (each costs ~0.1 s)
        s = int(g["seeds"][int(r["chosen"][row])])
        x, it = rwr_oracle.diffusion_profile(m0, s, sp_idx, sinks, synth.RWR_ALPHA, synth.RWR_TOL, synth.RWR_MAX_ITER)
        assert it == int(r["iters"][row])
        assert np.array_equal(x[r["sample_idx"]], r["profiles_sampled"][row])
        assert np.array_equal(x[r["top_idx"][row]], r["top_val"][row])
        if row < r["profiles_full"].shape[0]:
            assert np.array_equal(x, r["profiles_full"][row])


def test_rank_oracle_matches_reference_golden():
    g = load_golden("msi_covid_graph.npz")
    r = load_golden("rank_covid.npz")
    names = [str(x) for x in g["names"]]
    embs = synth.synth_embeddings(len(names), int(r["d"]), seed=int(r["emb_seed"])).astype(np.float64)
    is_drug = g["node_type"] == synth.DRUG
    rows, prox = rank_oracle.drug_proximity_ranking(names, embs, is_drug)
    assert np.array_equal(rows, r["drug_ranked_rows"])
    np.testing.assert_allclose(prox, r["drug_proximities_ranked"], rtol=1e-13)
    np.testing.assert_allclose(rank_oracle.protein_proximity_table(names, embs, r["protein_mask"]),
                               r["protein_proximities"], rtol=1e-12, atol=1e-15)
    d_rows, d_vals = rank_oracle.diffusion_drug_ranking(r["covid_profile"], is_drug)
    assert np.array_equal(d_vals, r["diffusion_values_ranked"])
    clear = np.concatenate([[True], np.diff(d_vals) != 0]) & np.concatenate([np.diff(d_vals) != 0, [True]])
    assert np.array_equal(d_rows[clear], r["diffusion_ranked_rows"][clear])          # ties excepted


def test_golden_msi_object_is_a_real_digraph():
    import networkx as nx
    msi = GoldenMSI("msi_covid_graph.npz", "rank_covid.npz")
    assert isinstance(msi.graph, nx.DiGraph) and msi.graph.number_of_edges() == 132392
    assert "NodeCovid" in msi.indications_in_graph and msi.graph.nodes["NodeCovid"]["type"] == "indication"
    assert len(msi.drug_or_indication2proteins["NodeCovid"]) > 0


def test_gcn_oracle_on_knn2000_golden():
    g = load_golden("gcn_knn2000.npz")
    n, d = g["embs"].shape
    assert n == 2000
    adj = torch.sparse_coo_tensor(torch.from_numpy(g["coo_indices"]), torch.from_numpy(g["coo_values"]), (n, n))
    tr = gcn_oracle.Trainer(g["embs"], adj, d, num_layers=2, layer_decay=0.3, lr=float(g["lr"]), w_init=g["w_init"])
    offs = np.concatenate([[0], np.cumsum(g["batch_sizes"])])
    losses = []
    for it in range(len(g["batch_sizes"])):
        loss, h, grads = tr.step(torch.from_numpy(g["batches_flat"][offs[it]:offs[it + 1]]), float(g["beta"]))
        losses.append(loss.item())
    assert_close(np.array(losses), g["losses"], what="losses")
    assert_close(h.numpy(), g["h_last"], what="hidden_emb last")


def test_gcn_oracle_on_real_msi_edgelist_golden():
    """One Adam step of train.py:155-184 on the real no-PPI MSI weighted edgelist (asymmetric Â)."""
    gr = load_golden("msi_noppi_graph.npz")
    g = load_golden("gcn_msi_noppi_edge.npz")
    n = len(gr["names"])
    a = sp.csr_matrix((gr["weight"], (gr["src"], gr["dst"])), shape=(n, n))
    a_hat = gcn_oracle.preprocess_graph(a)
    a_hat.sort_indices()
    assert np.array_equal(a_hat.indptr, g["ahat_indptr"]) and np.array_equal(a_hat.indices, g["ahat_indices"])
    assert np.array_equal(a_hat.data.astype(np.float32), g["ahat_data_f32"])
    assert (abs(a_hat - a_hat.T) > 1e-12).nnz > 0                  # genuinely asymmetric: backward needs Âᵀ
    embs = synth.synth_embeddings(n, int(g["d"]), seed=int(g["seed"]))
    tr = gcn_oracle.Trainer(embs, gcn_oracle.to_sparse_tensor(a_hat), int(g["d"]), lr=float(g["lr"]),
                            w_init=g["w_init"])
    b0 = torch.from_numpy(g["batches_flat"][:int(g["batch_sizes"][0])])
    loss, h, grads = tr.step(b0, float(g["beta"]))
    assert_close(loss.item(), g["losses"][0], what="loss step0")
    assert_close(h.numpy()[g["rows_kept"]], g["h_step0_rows"], what="hidden_emb rows step0")
    for nm, gr_ in zip(("dw1", "db1", "dw2", "db2"), grads):
        assert_close(gr_.numpy(), g[f"{nm}_step0"], what=f"{nm} step0")
