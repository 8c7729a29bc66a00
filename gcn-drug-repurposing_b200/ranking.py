"""The consumer step right after the path, on the device:

    drug_proximity_ranking      predict_drug.py:49-74   normalize(GCN embs) -> drug_embs @ covid_emb -> argsort[::-1]
    protein_proximity_table     run_covid.py:294-306    the same proximities for every named protein
    diffusion_drug_ranking      predict_drug.py:111-124 rank the drug entries of one diffusion profile

Same inputs (node names in .embs.txt row order, the fp64 embedding matrix np.loadtxt returns, the MSI object) and
the same return values as the reference code; the normalisation, the dot products and the sort run in libgdr
(csrc/rank.cu, fp64).  Ties in the scores are ordered by ascending position here and in an unspecified order by
numpy's quicksort: parity on ranks excepts ties.
"""
import numpy as np
import torch

from . import _ffi


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("gcn-drug-repurposing_b200 needs a CUDA device: there is no CPU implementation")
    return torch.device("cuda", torch.cuda.current_device())


def rank_by_embedding(node_embs, query_index, select_index=None, normalize=True):
    """node_embs: [n, d] array / tensor (fp64 on the device after conversion).  Returns
    (order int64[m] into select_index by descending proximity, proximities fp64[m] aligned with select_index)."""
    L = _ffi.lib()
    dev = _device()
    e = torch.as_tensor(np.asarray(node_embs, dtype=np.float64) if not isinstance(node_embs, torch.Tensor)
                        else node_embs).to(device=dev, dtype=torch.float64).contiguous()
    n, d = e.shape
    sel = None
    m = n
    if select_index is not None:
        sel = torch.as_tensor(np.asarray(select_index, dtype=np.int32)).to(dev)
        m = int(sel.numel())
        if m and (int(sel.min()) < 0 or int(sel.max()) >= n):
            raise IndexError("rank_by_embedding: select_index outside the embedding rows")
    prox = torch.empty(max(m, 1), dtype=torch.float64, device=dev)[:m]
    order = torch.empty(max(m, 1), dtype=torch.int32, device=dev)[:m]
    ws = _ffi.workspace(L.gdr_rank_workspace_bytes(m), dev)
    _ffi.check(L.gdr_rank_embedding_f64(_ffi.ptr(e), n, d, 1 if normalize else 0, int(query_index), _ffi.ptr(sel), m,
                                        _ffi.ptr(prox), _ffi.ptr(order), None, _ffi.ptr(ws), ws.numel(),
                                        _ffi.stream()), "rank_embedding_f64")
    return order.cpu().numpy().astype(np.int64), prox.cpu().numpy()


def rank_values(values, select_index=None):
    """np.argsort(values[select_index])[::-1] on the device.  Returns (order int64[m], values[select_index])."""
    L = _ffi.lib()
    dev = _device()
    v = torch.as_tensor(np.asarray(values, dtype=np.float64)).to(dev).contiguous()
    sel = None
    m = int(v.numel())
    if select_index is not None:
        sel = torch.as_tensor(np.asarray(select_index, dtype=np.int32)).to(dev)
        m = int(sel.numel())
        if m and (int(sel.min()) < 0 or int(sel.max()) >= v.numel()):
            raise IndexError("rank_values: select_index outside the value vector")
    out = torch.empty(max(m, 1), dtype=torch.float64, device=dev)[:m]
    order = torch.empty(max(m, 1), dtype=torch.int32, device=dev)[:m]
    ws = _ffi.workspace(L.gdr_rank_workspace_bytes(m), dev)
    _ffi.check(L.gdr_rank_values_f64(_ffi.ptr(v), _ffi.ptr(sel), m, _ffi.ptr(out), _ffi.ptr(order), None,
                                     _ffi.ptr(ws), ws.numel(), _ffi.stream()), "rank_values_f64")
    return order.cpu().numpy().astype(np.int64), out.cpu().numpy()


def _display_names(msi, nodes):
    # predict_drug.py:71-72: `drug if msi.node2name[drug] is np.nan else msi.node2name[drug]`
    return [n if msi.node2name[n] is np.nan else msi.node2name[n] for n in nodes]


def drug_proximity_ranking(node_names, node_embs, msi, query='NodeCovid', gcn=True):
    """predict_drug.py:49-74.  node_names: first column of the node2vec .embs.txt (row order of node_embs);
    node_embs: np.loadtxt(gcn_emb_file) (gcn=True: L2-normalised first, :52-53) or the node2vec vectors.
    Returns (drugs_name_ranked, drugs_ranked, proximities[ranked])."""
    node_names = list(node_names)
    q = node_names.index(query)
    drug_rows = [i for i, node in enumerate(node_names) if msi.graph.nodes[node]['type'] == 'drug']
    drug_names = [node_names[i] for i in drug_rows]
    order, prox = rank_by_embedding(node_embs, q, drug_rows, normalize=gcn)
    drugs_ranked = [drug_names[i] for i in order]
    return _display_names(msi, drugs_ranked), drugs_ranked, prox[order]


def protein_proximity_table(nodes, gcn_embs, msi, query='NodeCovid'):
    """run_covid.py:294-306: proximity of every protein that has a display name to the query node, in node order.
    Returns (protein_names, protein_proximities list)."""
    nodes = list(nodes)
    q = nodes.index(query)
    rows = [i for i, node in enumerate(nodes)
            if msi.graph.nodes[node]['type'] == 'protein' and msi.node2name[node] is not np.nan]
    _order, prox = rank_by_embedding(gcn_embs, q, rows, normalize=True)
    return [msi.node2name[nodes[i]] for i in rows], list(prox)


def diffusion_drug_ranking(profile, msi):
    """predict_drug.py:111-124: rank the drugs by their entry in one diffusion profile (msi.nodelist order).
    Returns (drugs_name_ranked, drugs_ranked, proximities[ranked])."""
    assert len(profile) == len(msi.nodelist)
    rows = [i for i, node in enumerate(msi.nodelist) if msi.graph.nodes[node]['type'] == 'drug']
    drugs = [msi.nodelist[i] for i in rows]
    order, vals = rank_values(profile, rows)
    drugs_ranked = [drugs[i] for i in order]
    return _display_names(msi, drugs_ranked), drugs_ranked, vals[order]
