"""Deterministic synthetic inputs of the shapes BASELINE.json names (the real blobs —
whole_graph*.weighted.edgelist, *.embs.txt, data/protein_to_protein.tsv — are listed in the
reference's .MISSING_LARGE_BLOBS and /root/reference does not exist on the GPU box).

  msi_synth        a multiscale-interactome-shaped graph with the published dimensions
                   (29,959 nodes / 478,728 undirected edges: 1,661 drugs, 840 indications,
                   17,660 proteins, 9,798 biological functions) and the reference's typed edge
                   weights (multiscale/msi/msi.py:230-262 restated on index arrays)
  synth_embeddings node2vec-like fp32 embeddings (rows not unit-norm) + the two .embs.txt formats
  rmat_coo         Graph500 R-MAT edge generator (torch, any device)
"""
import numpy as np

# node-type codes (order = nodelist order: drugs, indications, proteins, functions)
DRUG, INDICATION, PROTEIN, FUNCTION = 0, 1, 2, 3

MSI_COUNTS = dict(drugs=1661, indications=840, proteins=17660, functions=9798,
                  drug_protein=8568, indication_protein=25212, protein_protein=387626,
                  protein_function=34777, function_function=22545)

# predict_drug.py:85-92
MSI_WEIGHTS = {
    'down_functional_pathway': 4.4863053901688685,
    'indication': 3.541889556309463,
    'functional_pathway': 6.583155399238509,
    'up_functional_pathway': 2.09685000906964,
    'protein': 4.396695660380823,
    'drug': 3.2071696595616364,
}
RWR_ALPHA = 0.8595436247434408   # predict_drug.py:82
RWR_TOL = 1e-06
RWR_MAX_ITER = 1000


def _unique_pairs(rng, n_pairs, draw, symmetric, max_rounds=64):
    """Draw until n_pairs distinct (a,b) pairs exist (a != b when symmetric; unordered when symmetric)."""
    seen = np.empty((0, 2), dtype=np.int64)
    for _ in range(max_rounds):
        need = n_pairs - len(seen)
        if need <= 0:
            break
        a, b = draw(int(need * 1.3) + 16)
        if symmetric:
            keep = a != b
            a, b = a[keep], b[keep]
            lo, hi = np.minimum(a, b), np.maximum(a, b)
            a, b = lo, hi
        cand = np.concatenate([seen, np.stack([a, b], 1)])
        _, first = np.unique(cand, axis=0, return_index=True)
        seen = cand[np.sort(first)]
    if len(seen) < n_pairs:
        raise RuntimeError("msi_synth: could not draw enough distinct edges")
    return seen[:n_pairs]


def msi_synth(seed=20200430, scale=1.0):
    """Returns a dict describing the weighted, directed MSI-shaped graph:
      n, node_type[n], src[E2], dst[E2], weight[E2] (E2 = 2 * undirected edges; every undirected
      edge appears in both directions with its own per-source typed weight), names[n],
      seeds (drug + indication node ids, nodelist order), seed_proteins {seed -> sorted protein ids}.
    `scale` shrinks every count proportionally (tests use scale≈0.01)."""
    c = {k: max(int(round(v * scale)), 4) for k, v in MSI_COUNTS.items()}
    rng = np.random.default_rng(seed)
    nd, ni, npr, nf = c['drugs'], c['indications'], c['proteins'], c['functions']
    off_d, off_i, off_p, off_f = 0, nd, nd + ni, nd + ni + npr
    n = nd + ni + npr + nf
    node_type = np.concatenate([np.full(nd, DRUG), np.full(ni, INDICATION), np.full(npr, PROTEIN),
                                np.full(nf, FUNCTION)]).astype(np.int8)

    # degree-preferential protein popularity (Chung-Lu style, exponent ~2.5)
    pop = (np.arange(1, npr + 1, dtype=np.float64)) ** (-1.0 / 1.5)
    pop = pop[rng.permutation(npr)]
    pop /= pop.sum()
    fpop = (np.arange(1, nf + 1, dtype=np.float64)) ** (-0.5)
    fpop = fpop[rng.permutation(nf)]
    fpop /= fpop.sum()

    def prot(k):
        return rng.choice(npr, size=k, p=pop)

    max_dp = min(c['drug_protein'], nd * npr // 2)
    dp = _unique_pairs(rng, max_dp, lambda k: (rng.integers(0, nd, k), prot(k)), False)
    ip = _unique_pairs(rng, min(c['indication_protein'], ni * npr // 2),
                       lambda k: (rng.integers(0, ni, k), prot(k)), False)
    pp = _unique_pairs(rng, min(c['protein_protein'], npr * (npr - 1) // 4), lambda k: (prot(k), prot(k)), True)
    pf = _unique_pairs(rng, min(c['protein_function'], npr * nf // 2),
                       lambda k: (prot(k), rng.choice(nf, size=k, p=fpop)), False)
    # function hierarchy: child -> parent with child > parent (a DAG, so no pair has both directions)
    ff = _unique_pairs(rng, min(c['function_function'], nf * (nf - 1) // 4),
                       lambda k: (rng.integers(0, nf, k), rng.integers(0, nf, k)), True)
    # every drug / indication gets at least one protein (they are the seeds of the diffusion)
    miss_d = np.setdiff1d(np.arange(nd), dp[:, 0])
    miss_i = np.setdiff1d(np.arange(ni), ip[:, 0])
    if len(miss_d):
        dp[-len(miss_d):, 0] = miss_d   # re-point the last drawn edges (keeps the edge count)
        dp = np.unique(dp, axis=0)
    if len(miss_i):
        ip[-len(miss_i):, 0] = miss_i
        ip = np.unique(ip, axis=0)

    # group code of a directed edge u->v = which entry of MSI_WEIGHTS applies (msi.py:230-253):
    # successor type, except function->function which splits into up (v is a successor of u in the
    # hierarchy digraph) and down (v is a predecessor).
    G_DRUG, G_IND, G_PROT, G_FUNC, G_UP, G_DOWN = range(6)
    gw = np.array([MSI_WEIGHTS['drug'], MSI_WEIGHTS['indication'], MSI_WEIGHTS['protein'],
                   MSI_WEIGHTS['functional_pathway'], MSI_WEIGHTS['up_functional_pathway'],
                   MSI_WEIGHTS['down_functional_pathway']])
    src, dst, grp = [], [], []

    def add(u, v, g_uv, g_vu):
        src.extend([u, v]), dst.extend([v, u]), grp.extend([np.full(len(u), g_uv), np.full(len(u), g_vu)])

    add(dp[:, 0] + off_d, dp[:, 1] + off_p, G_PROT, G_DRUG)
    add(ip[:, 0] + off_i, ip[:, 1] + off_p, G_PROT, G_IND)
    add(pp[:, 0] + off_p, pp[:, 1] + off_p, G_PROT, G_PROT)
    add(pf[:, 0] + off_p, pf[:, 1] + off_f, G_FUNC, G_PROT)
    add(ff[:, 1] + off_f, ff[:, 0] + off_f, G_UP, G_DOWN)   # child(hi id) -> parent(lo id) is "up"
    src = np.concatenate(src).astype(np.int64)
    dst = np.concatenate(dst).astype(np.int64)
    grp = np.concatenate(grp).astype(np.int64)
    # weight = weights[group] / (number of successors of src in that group)   (msi.py:255-262)
    key = src * 6 + grp
    cnt = np.bincount(key, minlength=n * 6)
    weight = gw[grp] / cnt[key].astype(np.float64)

    order = np.lexsort((dst, src))
    src, dst, weight = src[order], dst[order], weight[order]
    names = ([f"DB{i:05d}" for i in range(nd)] + [f"C{i:07d}" for i in range(ni)] +
             [str(1000 + i) for i in range(npr)] + [f"GO:{i:07d}" for i in range(nf)])
    seeds = np.concatenate([np.arange(nd) + off_d, np.arange(ni) + off_i]).astype(np.int64)
    seed_proteins = {}
    for tbl, off in ((dp, off_d), (ip, off_i)):
        o = np.argsort(tbl[:, 0], kind='stable')
        t = tbl[o]
        bounds = np.searchsorted(t[:, 0], np.arange(t[:, 0].max() + 2))
        for s in range(len(bounds) - 1):
            if bounds[s + 1] > bounds[s]:
                seed_proteins[int(s + off)] = np.sort(t[bounds[s]:bounds[s + 1], 1] + off_p)
    return dict(n=n, node_type=node_type, src=src, dst=dst, weight=weight, names=names, seeds=seeds,
                seed_proteins=seed_proteins, n_undirected=len(src) // 2)


def synth_embeddings(n, d, seed=1):
    """node2vec-like rows: Gaussian direction, per-node norm ~ U(0.5, 3), fp32."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, d))
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    x *= rng.uniform(0.5, 3.0, size=(n, 1))
    return x.astype(np.float32)


def write_node2vec_embs(path, names, embs):
    """multiscale/openne/node2vec.py:40-47 format: header 'N dim', then 'node v1 ... vd'."""
    with open(path, 'w') as f:
        f.write(f"{embs.shape[0]} {embs.shape[1]}\n")
        for name, row in zip(names, embs):
            f.write(f"{name} {' '.join(str(x) for x in row)}\n")


def read_node2vec_embs(path):
    """train.py:79-80 equivalent (skip header, drop the name column) -> (names, fp64 [N, d])."""
    names, rows = [], []
    with open(path) as f:
        next(f)
        for line in f:
            parts = line.split()
            names.append(parts[0])
            rows.append(np.array(parts[1:], dtype=np.float64))
    return names, np.stack(rows)


def rmat_coo(scale, edge_factor, seed, device, a=0.57, b=0.19, c=0.19, chunk=1 << 26):
    """Graph500 R-MAT: 2^scale vertices, edge_factor*2^scale directed edges (duplicates and
    self-loops kept).  Returns int64 (row, col) torch tensors on `device`."""
    import torch
    n_edges = edge_factor << scale
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    rows, cols = [], []
    ab, c_norm, a_norm = a + b, c / (1.0 - (a + b)), a / (a + b)
    for start in range(0, n_edges, chunk):
        m = min(chunk, n_edges - start)
        r = torch.zeros(m, dtype=torch.int64, device=device)
        cc = torch.zeros(m, dtype=torch.int64, device=device)
        for _ in range(scale):
            ii = torch.rand(m, generator=g, device=device) > ab
            jj = torch.rand(m, generator=g, device=device) > torch.where(ii, c_norm, a_norm)
            r = (r << 1) | ii.to(torch.int64)
            cc = (cc << 1) | jj.to(torch.int64)
        rows.append(r), cols.append(cc)
    return torch.cat(rows), torch.cat(cols)
