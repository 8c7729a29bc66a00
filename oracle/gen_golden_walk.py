"""Generates tests/golden/walk_small.npz from the REFERENCE's multiscale/openne/walker.py (pure Python, imported from
/root/reference): a small weighted DiGraph, the per-node and per-edge alias tables the reference builds, converted to
the probabilities they sample, and asserts that oracle/walk_oracle.py gives the same numbers (to the float32 rounding
of the tables' q arrays).

    python oracle/gen_golden_walk.py
"""
import importlib.util
import os
import sys
import types

import networkx as nx
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import walk_oracle  # noqa: E402

REF = "/root/reference/multiscale/openne/walker.py"


def main():
    spec = importlib.util.spec_from_file_location("ref_walker", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    rng = np.random.default_rng(5)
    n = 40
    g = nx.DiGraph()
    g.add_nodes_from(str(i) for i in range(n))
    for _ in range(260):
        u, v = rng.integers(0, n, 2)
        if u != v:
            w = float(rng.uniform(0.2, 3.0))
            g.add_edge(str(u), str(v), weight=w)
            if rng.random() < 0.6:
                g.add_edge(str(v), str(u), weight=float(rng.uniform(0.2, 3.0)))
    g.remove_edges_from(list(g.out_edges("7")))               # a dead end: walks stop there
    wrap = types.SimpleNamespace(G=g, node_size=n, look_up_dict={v: i for i, v in enumerate(g.nodes())})
    p, q = 0.5, 2.0
    w = ref.Walker(wrap, p, q, 1)
    w.preprocess_transition_probs()
    nodes = list(g.nodes())
    idx = {v: i for i, v in enumerate(nodes)}
    e_src, e_dst, e_w = [], [], []
    for u, v, d in g.edges(data=True):
        e_src.append(idx[u]); e_dst.append(idx[v]); e_w.append(d["weight"])
    # second-order distributions, flattened: for edge k, entries [off[k], off[k+1]) over the neighbours of dst
    off, nb, pr = [0], [], []
    for (u, v), (J, qq) in w.alias_edges.items():
        probs = walk_oracle.alias_to_probs(J, qq)
        nbrs, exp = walk_oracle.second_order_probs(g, p, q, u, v)
        assert np.allclose(probs, exp, atol=2e-6), (u, v, probs, exp)
        nb += [idx[x] for x in nbrs]
        pr += list(exp)
        off.append(len(nb))
    keys = np.array([[idx[u], idx[v]] for (u, v) in w.alias_edges.keys()], dtype=np.int64)
    f_off, f_nb, f_pr = [0], [], []
    for v in nodes:
        J, qq = w.alias_nodes[v]
        if len(J):
            nbrs, exp = walk_oracle.first_order_probs(g, v)
            assert np.allclose(walk_oracle.alias_to_probs(J, qq), exp, atol=2e-6)
            f_nb += [idx[x] for x in nbrs]
            f_pr += list(exp)
        f_off.append(len(f_nb))
    path = os.path.join(ROOT, "tests", "golden", "walk_small.npz")
    np.savez_compressed(path, n=n, p=p, q=q, edge_src=np.array(e_src), edge_dst=np.array(e_dst), edge_w=np.array(e_w),
                        so_keys=keys, so_off=np.array(off), so_nbr=np.array(nb), so_prob=np.array(pr),
                        fo_off=np.array(f_off), fo_nbr=np.array(f_nb), fo_prob=np.array(f_pr))
    print("wrote", path, "edges", len(e_src), "second-order tables", len(keys))


if __name__ == "__main__":
    main()
