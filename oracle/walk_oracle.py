"""TEST INFRASTRUCTURE ONLY.  Exact transition probabilities of the reference's walk generators
(multiscale/openne/walker.py), restated without the alias tables; pinned against the tables the reference itself
builds by oracle/gen_golden_walk.py (tests/golden/walk_small.npz)."""
import numpy as np


def first_order_probs(G, node):
    """walker.py:133-140: weight / sum of weights over G.neighbors(node), in neighbour order."""
    nbrs = list(G.neighbors(node))
    w = np.array([G[node][x]['weight'] for x in nbrs], dtype=np.float64)
    return nbrs, w / w.sum()


def second_order_probs(G, p, q, src, dst):
    """walker.py:106-125 (get_alias_edge): the walk came src -> dst; distribution of the next node."""
    nbrs = list(G.neighbors(dst))
    u = []
    for x in nbrs:
        if x == src:
            u.append(G[dst][x]['weight'] / p)
        elif G.has_edge(x, src):
            u.append(G[dst][x]['weight'])
        else:
            u.append(G[dst][x]['weight'] / q)
    u = np.array(u, dtype=np.float64)
    return nbrs, u / u.sum()


def alias_to_probs(J, q):
    """The distribution an alias table (walker.py:164-207) samples: column k is hit with q[k]/K directly and with
    (1 - q[j])/K through every j whose alias is k."""
    K = len(J)
    pr = np.asarray(q, dtype=np.float64) / K
    for j in range(K):
        pr[J[j]] += (1.0 - float(q[j])) / K
    return pr
