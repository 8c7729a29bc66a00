"""ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the consumer ranking step that follows the path
(SURVEY §8f-2).  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.

Parity status: PINNED AGAINST THE REFERENCE ITSELF — oracle/gen_golden_msi.py executes the reference's own source
lines (predict_drug.py:48-74, :111-124 and run_covid.py:293-306, read from /root/reference at generation time) on
the covid-variant MSI and asserts equal rankings and bit-equal proximities before writing tests/golden/rank_covid.npz.
"""
import numpy as np
from sklearn.preprocessing import normalize


def drug_proximity_ranking(node_names, node_embs, is_drug, query='NodeCovid', gcn=True):
    """predict_drug.py:49-74.  Returns (ranked row indices of the drugs, proximities in that order)."""
    node_names = list(node_names)
    node_embs = np.asarray(node_embs, dtype=np.float64)
    if gcn:
        node_embs = normalize(node_embs, axis=1)                     # :53
    covid_emb = np.array(node_embs[node_names.index(query)])        # :57
    rows = [i for i in range(len(node_names)) if is_drug[i]]
    drug_embs = np.array([node_embs[i] for i in rows])               # :59-65
    proximities = np.matmul(drug_embs, covid_emb)                    # :66
    ranked = np.argsort(np.array(proximities))[::-1]                 # :70
    return np.asarray(rows)[ranked], proximities[ranked]


def protein_proximity_table(nodes, gcn_embs, protein_mask, query='NodeCovid'):
    """run_covid.py:294-306.  protein_mask[i]: node i is a protein with a display name."""
    gcn_embs = normalize(np.asarray(gcn_embs, dtype=np.float64), axis=1)
    covid_emb = gcn_embs[list(nodes).index(query)]
    protein_embs = gcn_embs[np.asarray(protein_mask, dtype=bool)]
    return np.matmul(protein_embs, covid_emb)


def diffusion_drug_ranking(profile, is_drug):
    """predict_drug.py:111-124.  Returns (ranked positions among the drugs, their profile entries in that order)."""
    rows = [i for i in range(len(profile)) if is_drug[i]]
    proximities = np.array([profile[i] for i in rows])
    ranked = np.argsort(proximities)[::-1]
    return np.asarray(rows)[ranked], proximities[ranked]
