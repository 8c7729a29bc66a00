"""Golden fixtures on the REAL multiscale interactome (VERDICT r1 "real-data fixtures").

Builds the no-PPI MSI (22,550 nodes / 182,204 directed edges; data/protein_to_protein.tsv is in the reference's
.MISSING_LARGE_BLOBS) and the covid variant (indication2protein = data/covid_to_protein.tsv; 20,228 nodes,
NodeCovid the only indication) with the reference's OWN code (multiscale/msi/msi.py:207-262 from
/root/reference/data/*.tsv), weights them (msi.py:255-262, weight dict of predict_drug.py:85-92), runs the
reference's DiffusionProfiles seed by seed (diffusion_profiles.py:103-110) on >= 64 seeds of each, runs the
reference GCN (modules/model.py, helpers/helper.py) for 3 Adam steps on the no-PPI weighted edgelist pushed through
preprocess_graph (SURVEY §0.3 path B), pins the oracle restatement bit for bit on all of it, and writes

    tests/golden/msi_noppi_graph.npz, msi_covid_graph.npz     node names/types, directed weighted edges, seeds
    tests/golden/rwr_msi_noppi.npz,  rwr_msi_covid.npz        iteration counts + profiles (see below)
    tests/golden/gcn_msi_noppi_edge.npz                       losses, beta, grads, sampled hidden_emb rows
    tests/golden/rank_covid.npz                               drug / protein rankings of the consumer step, produced
                                                              by exec'ing predict_drug.py:48-74,111-121 and
                                                              run_covid.py:293-306 from the reference tree

The TSVs do not travel to the GPU box, hence the committed graph files.  Profiles are 8*N bytes per seed, so the
fixture keeps 12 of them whole and, for every seed, 512 fixed sampled entries plus the 64 largest entries.

    python oracle/gen_golden_msi.py          # run in the build container; exits non-zero on any mismatch

TEST INFRASTRUCTURE ONLY.  Nothing here is imported by the product package.
"""
import os
import sys
import warnings

import numpy as np
import scipy.sparse as sp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
warnings.filterwarnings("ignore")

from oracle import gcn_oracle, rwr_oracle  # noqa: E402
from oracle.gen_golden import import_reference, same, csr_parts  # noqa: E402
import importlib  # noqa: E402

synth = importlib.import_module("gcn-drug-repurposing_b200.synth")
GOLD = os.path.join(ROOT, "tests", "golden")
N_SEEDS, N_FULL, N_SAMPLE, N_TOP = 64, 12, 512, 64
TYPE_CODE = {"drug": synth.DRUG, "indication": synth.INDICATION, "protein": synth.PROTEIN,
             "functional_pathway": synth.FUNCTION}


def build_msi(covid):
    from msi.msi import MSI, DRUG_PROTEIN, INDICATION_PROTEIN, PROTEIN_FUNCTIONAL_PATHWAY, \
        FUNCTIONAL_PATHWAY_FUNCTIONAL_PATHWAY
    cwd = os.getcwd()
    os.chdir(REF)                                   # TSV paths are cwd-relative (msi.py:35-43)
    try:
        kw = dict(edges=[DRUG_PROTEIN, INDICATION_PROTEIN, PROTEIN_FUNCTIONAL_PATHWAY,
                         FUNCTIONAL_PATHWAY_FUNCTIONAL_PATHWAY])          # no PPI layer: blob absent upstream
        if covid:
            kw["indication2protein_file_path"] = "data/covid_to_protein.tsv"
        msi = MSI(**kw)
        msi.load()
    finally:
        os.chdir(cwd)
    # the reference iterates sets of strings here (msi.py:180,183): pin the order so the fixture is reproducible
    msi.drugs_in_graph = sorted(msi.drugs_in_graph)
    msi.indications_in_graph = sorted(msi.indications_in_graph)
    msi.weight_graph(synth.MSI_WEIGHTS)
    return msi


def graph_arrays(msi):
    names = list(msi.nodelist)
    idx = msi.node2idx
    src, dst, w = [], [], []
    for u, v, wt in msi.graph.edges(data="weight"):
        src.append(idx[u]), dst.append(idx[v]), w.append(wt)
    src, dst, w = np.asarray(src, np.int32), np.asarray(dst, np.int32), np.asarray(w, np.float64)
    seeds = [idx[t] for t in msi.drugs_in_graph + msi.indications_in_graph]
    ptr, prot = [0], []
    for t in msi.drugs_in_graph + msi.indications_in_graph:
        prot.extend(sorted(idx[p] for p in msi.drug_or_indication2proteins[t]))
        ptr.append(len(prot))
    return dict(names=np.array(names), node_type=np.array([TYPE_CODE[msi.node2type[n]] for n in names], np.int8),
                src=src, dst=dst, weight=w, seeds=np.asarray(seeds, np.int32),
                n_drugs=np.int64(len(msi.drugs_in_graph)),
                seed_protein_ptr=np.asarray(ptr, np.int64), seed_protein_idx=np.asarray(prot, np.int32))


def rwr_case(name, msi, DiffusionProfiles, counters, ga):
    dp = DiffusionProfiles(alpha=synth.RWR_ALPHA, max_iter=synth.RWR_MAX_ITER, tol=synth.RWR_TOL,
                           weights=None, num_cores=1, save_load_file_path=None)
    dp.get_initial_M(msi)
    n = len(msi.nodelist)
    m0 = rwr_oracle.initial_matrix(n, ga["src"], ga["dst"], ga["weight"])
    assert (dp.initial_M != m0).nnz == 0, f"{name}: oracle initial_matrix != reference get_initial_M"
    all_seeds = msi.drugs_in_graph + msi.indications_in_graph
    rng = np.random.default_rng(7)
    chosen = np.sort(rng.choice(len(all_seeds), size=min(N_SEEDS, len(all_seeds)), replace=False))
    if "NodeCovid" in all_seeds and all_seeds.index("NodeCovid") not in chosen:
        chosen[-1] = all_seeds.index("NodeCovid")           # the seed predict_drug.py:112 ranks the drugs by
        chosen = np.sort(chosen)
    sample_idx = np.sort(rng.choice(n, size=N_SAMPLE, replace=False)).astype(np.int32)
    seed_proteins = {int(s): ga["seed_protein_idx"][ga["seed_protein_ptr"][i]:ga["seed_protein_ptr"][i + 1]]
                     for i, s in enumerate(ga["seeds"])}
    sinks = [int(s) for s in ga["seeds"]]
    iters, full, sampled, top_idx, top_val, sums = [], [], [], [], [], []
    for k, ci in enumerate(chosen):
        nm = all_seeds[ci]
        M = dp.convert_M_to_make_all_drugs_indications_sinks_except_selected(msi, [nm])
        M, S = dp.refine_M_S(M)
        per = dp.get_personalization_dictionary([nm], msi.nodelist)
        counters["absolute"] = 0
        x = dp.power_iteration(M, S, msi.nodelist, per)
        it = counters["absolute"]
        ox, oit = rwr_oracle.diffusion_profile(m0, msi.node2idx[nm], seed_proteins, sinks, synth.RWR_ALPHA,
                                               synth.RWR_TOL, synth.RWR_MAX_ITER)
        same(x, ox, f"{name}.profile[{nm}]")
        assert it == oit, (nm, it, oit)
        iters.append(it)
        if k < N_FULL:
            full.append(x)
        sampled.append(x[sample_idx])
        t = np.argsort(-x, kind="stable")[:N_TOP]
        top_idx.append(t.astype(np.int32)), top_val.append(x[t]), sums.append(x.sum())
    np.savez_compressed(os.path.join(GOLD, f"rwr_{name}.npz"), chosen=chosen.astype(np.int32),
                        seed_names=np.array([all_seeds[c] for c in chosen]), iters=np.asarray(iters, np.int32),
                        profiles_full=np.stack(full), sample_idx=sample_idx, profiles_sampled=np.stack(sampled),
                        top_idx=np.stack(top_idx), top_val=np.stack(top_val), sums=np.asarray(sums))
    print(f"rwr_{name}: n={n} nnz={m0.nnz} seeds={len(chosen)} iters {min(iters)}..{max(iters)}")


def gcn_edge_case(name, ref_model, ref_helper, ga, d=128, batch=2048, steps=3, seed=6, n_rows_kept=384):
    """train.py:99-184 on the weighted MSI edgelist as the adjacency (SURVEY §0.3 path B: asymmetric weights)."""
    n = len(ga["names"])
    x_adj = sp.csr_matrix((ga["weight"], (ga["src"], ga["dst"])), shape=(n, n))
    a_hat = ref_helper.preprocess_graph(x_adj)
    same(csr_parts(a_hat)[2], csr_parts(gcn_oracle.preprocess_graph(x_adj))[2], f"{name}.preprocess_graph")
    adj_t = ref_helper.convert_sparse_matrix_to_sparse_tensor(a_hat)
    o_adj_t = gcn_oracle.to_sparse_tensor(a_hat)
    embs = synth.synth_embeddings(n, d, seed=seed)
    torch.manual_seed(seed)
    np.random.seed(seed)
    model = ref_model.ResidualGraphConvolutionalNetwork(batch, n, num_layers=2, hidden_units=d, init_weights=1e-5,
                                                        layer_decay=0.3)
    np.random.seed(seed)
    trainer = gcn_oracle.Trainer(embs, o_adj_t, d, num_layers=2, layer_decay=0.3, lr=3e-4)
    same(model.gcn_layer.dense.weight.detach().numpy(), trainer.w1.detach().numpy(), f"{name}.init_w")
    w_init = model.gcn_layer.dense.weight.detach().numpy().copy()
    feats = torch.tensor(embs, dtype=torch.float32)
    opt = torch.optim.Adam(model.parameters(), lr=3e-4, weight_decay=0)
    loss_fcn = ref_model.GSS_loss(1.0).gss_loss
    g = torch.Generator().manual_seed(seed)
    perm = torch.randperm(n, generator=g)
    batches = [perm[i * batch:(i + 1) * batch] for i in range(steps)]
    rows_kept = np.sort(np.random.default_rng(seed).choice(n, size=n_rows_kept, replace=False))
    out = dict(seed=np.int64(seed), d=np.int64(d), lr=3e-4, layer_decay=0.3, alpha=1.0, beta_percentile=98.0,
               num_layers=np.int64(2), w_init=w_init, batches_flat=np.concatenate([b.numpy() for b in batches]),
               batch_sizes=np.array([len(b) for b in batches], np.int64), rows_kept=rows_kept,
               ahat_indptr=csr_parts(a_hat)[0], ahat_indices=csr_parts(a_hat)[1],
               ahat_data_f32=csr_parts(a_hat)[2].astype(np.float32))
    losses, beta = [], None
    for it, b in enumerate(batches):
        h = model(x=feats, adj=adj_t)
        if it == 0:
            h0 = h.cpu().data
            beta = np.percentile(np.dot(h0, h0.transpose(0, 1)).flatten(), 98.0)          # train.py:166-167
            out["h_step0_rows"] = h.detach().numpy()[rows_kept].copy()
        loss = loss_fcn(embs=h, beta=beta, index=b.detach())
        opt.zero_grad()
        loss.backward()
        grads = [p.grad.detach().numpy().copy() for p in model.parameters()]
        opt.step()
        o_loss, o_h, o_grads = trainer.step(b, beta)
        same(h.detach().numpy(), o_h.numpy(), f"{name}.step{it}.hidden_emb")
        same(loss.detach().numpy(), o_loss.numpy(), f"{name}.step{it}.loss")
        for a_, b_, w in zip(grads, o_grads, ("dw1", "db1", "dw2", "db2")):
            same(a_, b_.numpy(), f"{name}.step{it}.grad.{w}")
        losses.append(float(loss))
        if it in (0, steps - 1):
            tag = "step0" if it == 0 else "last"
            for w, g_ in zip(("dw1", "db1", "dw2", "db2"), grads):
                out[f"{w}_{tag}"] = g_
            out[f"h_{tag}_rows"] = h.detach().numpy()[rows_kept].copy()
    for w, p in zip(("w1", "b1", "w2", "b2"), model.parameters()):
        out[f"{w}_final"] = p.detach().numpy().copy()
    out["beta"] = np.float32(beta)
    out["losses"] = np.asarray(losses, np.float64)
    np.savez_compressed(os.path.join(GOLD, f"gcn_{name}.npz"), **out)
    print(f"gcn_{name}: n={n} d={d} nnz(Â)={a_hat.nnz} beta={beta:.6f} losses={losses}")


def _ref_lines(rel_path, first, last):
    """Source lines [first, last] (1-based, inclusive) of a reference file, dedented."""
    import textwrap
    with open(os.path.join(REF, rel_path)) as f:
        lines = f.readlines()
    return textwrap.dedent("".join(lines[first - 1:last]))


def rank_case(name, msi, ga, d=128, seed=9):
    """The consumer step executed from the reference's own source text (the scripts cannot be imported: they pull
    in packages that are not installed at import time) and pinned against oracle/rank_oracle.py."""
    import tempfile
    from sklearn.preprocessing import normalize
    from oracle import rank_oracle
    n = len(ga["names"])
    names = [str(x) for x in ga["names"]]
    embs = synth.synth_embeddings(n, d, seed=seed)                  # stands in for the GCN output (fp32, train.py:193)
    is_drug = ga["node_type"] == synth.DRUG
    with tempfile.TemporaryDirectory() as tmp:
        emb_file = os.path.join(tmp, "n2v.embs.txt")
        gcn_emb_file = os.path.join(tmp, "whole_graph_gcn_pathway.embs.txt")
        synth.write_node2vec_embs(emb_file, names, embs)
        np.savetxt(gcn_emb_file, embs)                              # train.py:193
        # ---- predict_drug.py:48-74 (gcn=True) -------------------------------------------------------------
        ns = dict(np=np, normalize=normalize, msi=msi, emb_file=emb_file, gcn=True, gcn_emb_file=gcn_emb_file)
        src = _ref_lines("predict_drug.py", 48, 73)
        exec(src, ns)
        ref_ranked = [names.index(x) for x in ns["drugs_ranked"]]
        ref_prox = ns["proximities"][ns["proximities_ranked_id"]]
        o_rows, o_prox = rank_oracle.drug_proximity_ranking(names, np.loadtxt(gcn_emb_file), is_drug)
        same(ref_ranked, o_rows, f"{name}.drug ranking")
        same(ref_prox, o_prox, f"{name}.drug proximities")
        # ---- run_covid.py:293-306 ------------------------------------------------------------------------------
        cwd = os.getcwd()
        os.chdir(tmp)
        try:
            ns2 = dict(np=np, normalize=normalize, msi=msi, emb_file=emb_file)
            exec(_ref_lines("run_covid.py", 293, 306), ns2)
        finally:
            os.chdir(cwd)
        pmask = np.array(ns2["protein_mask"], dtype=bool)
        same(np.array(ns2["protein_proximities"]), rank_oracle.protein_proximity_table(names, np.loadtxt(gcn_emb_file),
                                                                                        pmask), f"{name}.protein table")
    # ---- predict_drug.py:111-124 on the NodeCovid diffusion profile ----------------------------------------------
    n_ = len(names)
    m0 = rwr_oracle.initial_matrix(n_, ga["src"], ga["dst"], ga["weight"])
    seed_proteins = {int(s): ga["seed_protein_idx"][ga["seed_protein_ptr"][i]:ga["seed_protein_ptr"][i + 1]]
                     for i, s in enumerate(ga["seeds"])}
    covid = names.index("NodeCovid")
    profile, _ = rwr_oracle.diffusion_profile(m0, covid, seed_proteins, [int(s) for s in ga["seeds"]], synth.RWR_ALPHA,
                                              synth.RWR_TOL, synth.RWR_MAX_ITER)

    class _DP:
        drug_or_indication2diffusion_profile = {"NodeCovid": profile}
    ns3 = dict(np=np, msi=msi, dp_saved=_DP())
    exec(_ref_lines("predict_drug.py", 111, 121), ns3)
    d_rows = np.array([names.index(x) for x in ns3["drugs_ranked"]])
    d_vals = np.array(ns3["proximities"])[ns3["proximities_ranked_id"]]
    o_rows2, o_vals2 = rank_oracle.diffusion_drug_ranking(profile, is_drug)
    same(d_rows, o_rows2, f"{name}.diffusion ranking")
    same(d_vals, o_vals2, f"{name}.diffusion values")
    has_name = np.array([msi.node2name[nm] is not np.nan for nm in names])
    np.savez_compressed(os.path.join(GOLD, f"rank_{name}.npz"), emb_seed=np.int64(seed), d=np.int64(d),
                        drug_ranked_rows=np.asarray(ref_ranked, np.int32), drug_proximities_ranked=ref_prox,
                        protein_mask=pmask, protein_proximities=np.array(ns2["protein_proximities"]),
                        protein_names=np.array(ns2["protein_names"]), has_name=has_name,
                        display_name=np.array([str(msi.node2name[nm]) for nm in names]),
                        drugs_name_ranked_head=np.array([str(x) for x in ns["drugs_name_ranked"][:20]]),
                        covid_profile=profile, diffusion_ranked_rows=d_rows.astype(np.int32),
                        diffusion_values_ranked=d_vals)
    print(f"rank_{name}: {len(ref_ranked)} drugs ranked, {int(pmask.sum())} named proteins, top drug row {ref_ranked[0]}")


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref_model, ref_helper, DiffusionProfiles, counters = import_reference()
    for name, covid in (("msi_noppi", False), ("msi_covid", True)):
        msi = build_msi(covid)
        ga = graph_arrays(msi)
        print(f"{name}: {len(ga['names'])} nodes, {len(ga['src'])} directed edges, {len(ga['seeds'])} seeds "
              f"({int(ga['n_drugs'])} drugs)")
        np.savez_compressed(os.path.join(GOLD, f"{name}_graph.npz"), **ga)
        rwr_case(name, msi, DiffusionProfiles, counters, ga)
        if not covid:
            gcn_edge_case("msi_noppi_edge", ref_model, ref_helper, ga)
        else:
            rank_case("covid", msi, ga)
    print("real-MSI golden fixtures written; oracle == reference on every compared tensor")


if __name__ == "__main__":
    main()
