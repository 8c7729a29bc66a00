"""Rebuilds an MSI-like object (a REAL networkx.DiGraph + the attributes DiffusionProfiles and the ranking step
read, multiscale/msi/msi.py:148-205) from the committed real-interactome fixtures tests/golden/msi_*_graph.npz
(written by oracle/gen_golden_msi.py from /root/reference/data/*.tsv, which does not travel to the GPU box)."""
import numpy as np

from .conftest import load_golden

TYPE_NAME = {0: "drug", 1: "indication", 2: "protein", 3: "functional_pathway"}


class GoldenMSI:
    def __init__(self, graph_file, rank_file=None):
        import networkx as nx
        g = load_golden(graph_file)
        names = [str(x) for x in g["names"]]
        self.nodelist = names
        self.node2idx = {nm: i for i, nm in enumerate(names)}
        self.graph = nx.DiGraph()
        for nm, t in zip(names, g["node_type"]):
            self.graph.add_node(nm, type=TYPE_NAME[int(t)])
        self.graph.add_weighted_edges_from(
            (names[u], names[v], float(w)) for u, v, w in zip(g["src"], g["dst"], g["weight"]))
        nd = int(g["n_drugs"])
        seeds = [names[i] for i in g["seeds"]]
        self.drugs_in_graph, self.indications_in_graph = seeds[:nd], seeds[nd:]
        ptr, idx = g["seed_protein_ptr"], g["seed_protein_idx"]
        self.drug_or_indication2proteins = {s: set(names[p] for p in idx[ptr[i]:ptr[i + 1]])
                                            for i, s in enumerate(seeds)}
        self.arrays = g
        self.saved = []
        if rank_file is not None:
            r = load_golden(rank_file)
            self.node2name = {nm: (str(dn) if hn else np.nan)
                              for nm, dn, hn in zip(names, r["display_name"], r["has_name"])}

    # calculate_diffusion_profiles side effects (diffusion_profiles.py:125-130)
    def save_graph(self, path):
        self.saved.append("graph")

    def save_node2idx(self, path):
        self.saved.append("node2idx")

    def weight_graph(self, weights):
        self.saved.append("weighted")       # the fixture graph is already weighted (msi.py:255-262 ran upstream)

    def seed_proteins_by_index(self):
        g = self.arrays
        return {int(s): g["seed_protein_idx"][g["seed_protein_ptr"][i]:g["seed_protein_ptr"][i + 1]]
                for i, s in enumerate(g["seeds"])}
