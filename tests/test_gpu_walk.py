"""GPU tests of the walk generators (walker.py over csrc/walk.cu) against the transition probabilities of the
REFERENCE's alias tables (tests/golden/walk_small.npz, generated from multiscale/openne/walker.py by
oracle/gen_golden_walk.py).  The reference draws from numpy's global stream, the kernel from a counter-based
generator, so parity is distributional: every observed transition frequency must lie within 5 standard deviations of
the reference probability (and the structure — edges followed, dead ends, start order — must be exact)."""
import random
import types

import networkx as nx
import numpy as np
import pytest

from .conftest import load_golden

pytestmark = pytest.mark.gpu


def _graph(g):
    n = int(g["n"])
    nxg = nx.DiGraph()
    nxg.add_nodes_from(str(i) for i in range(n))
    for u, v, w in zip(g["edge_src"], g["edge_dst"], g["edge_w"]):
        nxg.add_edge(str(u), str(v), weight=float(w))
    return types.SimpleNamespace(G=nxg, node_size=n, look_up_dict={str(i): i for i in range(n)})


def _check_freq(counts, probs, what):
    tot = counts.sum()
    big = probs * tot >= 25                                     # normal approximation of the binomial
    if tot < 500 or not big.any():
        return 0
    sd = np.sqrt(np.maximum(probs * (1 - probs), 1e-12) / tot)
    dev = (np.abs(counts / tot - probs) / sd)[big]
    assert dev.max() < 5.0, f"{what}: frequency {counts / tot} vs probability {probs} ({dev.max():.1f} sigma, {tot} draws)"
    return 1


def test_node2vec_walks_follow_reference_transition_probabilities(gdr):
    g = load_golden("walk_small.npz")
    wrap = _graph(g)
    n = int(g["n"])
    w = gdr.walker.Walker(wrap, float(g["p"]), float(g["q"]), 1)
    w.preprocess_transition_probs()
    random.seed(3)
    walks, nodes = w.simulate_walks_array(num_walks=3000, walk_length=12)
    arr = walks.cpu().numpy()
    assert arr.shape == (3000 * n, 12) and nodes == [str(i) for i in range(n)]
    # structure: every pass starts once at every node; every step is an edge; -1 only after a dead end, then to the end
    for it in (0, 1, 2999):
        assert sorted(arr[it * n:(it + 1) * n, 0].tolist()) == list(range(n))
    adj = {(int(u), int(v)) for u, v in zip(g["edge_src"], g["edge_dst"])}
    outdeg = np.bincount(g["edge_src"], minlength=n)
    a, b = arr[:, :-1], arr[:, 1:]
    step = (a >= 0) & (b >= 0)
    pairs = set(zip(a[step].tolist(), b[step].tolist()))
    assert pairs <= adj, "a walk left the graph"
    ended = (a >= 0) & (b < 0)
    assert (outdeg[a[ended]] == 0).all(), "a walk stopped at a node that has out-neighbours"
    assert ((a < 0) <= (b < 0)).all()
    # first step: first-order distribution (walker.py:83-85)
    checked = 0
    for v in range(n):
        lo, hi = int(g["fo_off"][v]), int(g["fo_off"][v + 1])
        if hi == lo:
            continue
        nxt = arr[arr[:, 0] == v, 1]
        counts = np.array([(nxt == x).sum() for x in g["fo_nbr"][lo:hi]])
        checked += _check_freq(counts, g["fo_prob"][lo:hi], f"first step from {v}")
    assert checked >= n // 2
    # later steps: second-order distribution given (previous, current) (walker.py:86-92, 106-125)
    prev, cur, nxt = arr[:, :-2].ravel(), arr[:, 1:-1].ravel(), arr[:, 2:].ravel()
    ok = nxt >= 0
    key = prev[ok].astype(np.int64) * n + cur[ok]
    nx_ok = nxt[ok]
    order = np.argsort(key, kind="stable")
    key, nx_ok = key[order], nx_ok[order]
    checked = 0
    for k, (u, v) in enumerate(g["so_keys"]):
        lo, hi = int(g["so_off"][k]), int(g["so_off"][k + 1])
        if hi == lo:
            continue
        s, e = np.searchsorted(key, [u * n + v, u * n + v + 1])
        counts = np.array([(nx_ok[s:e] == x).sum() for x in g["so_nbr"][lo:hi]])
        assert counts.sum() == e - s
        checked += _check_freq(counts, g["so_prob"][lo:hi], f"step after {u}->{v}")
    assert checked >= len(g["so_keys"]) // 2, checked


def test_simulate_walks_list_form_and_seeding(gdr):
    """walker.py:93-106: a list of walks of node ids; `random.seed` fixes the start order and the device stream."""
    g = load_golden("walk_small.npz")
    wrap = _graph(g)
    w = gdr.walker.Walker(wrap, 0.5, 2.0, 1)
    w.preprocess_transition_probs()
    random.seed(11)
    a = w.simulate_walks(num_walks=2, walk_length=6)
    random.seed(11)
    b = w.simulate_walks(num_walks=2, walk_length=6)
    assert a == b and len(a) == 2 * int(g["n"])
    assert all(isinstance(x, str) for x in a[0]) and all(1 <= len(x) <= 6 for x in a)
    n = int(g["n"])
    order = [str(i) for i in range(n)]
    random.seed(11)                                             # a plain replay of walker.py:102: nothing but shuffles
    for it in range(2):
        random.shuffle(order)
        assert [x[0] for x in a[it * n:(it + 1) * n]] == order  # every pass starts in the reference's order
    assert ["7"] in a                                           # node 7 has no out-edges: a walk of length 1
    one = w.node2vec_walk(walk_length=5, start_node="3")
    assert one[0] == "3" and len(one) <= 5


def test_deepwalk_is_uniform_over_neighbours(gdr):
    g = load_golden("walk_small.npz")
    wrap = _graph(g)
    n = int(g["n"])
    bw = gdr.walker.BasicWalker(wrap, 1)
    random.seed(5)
    walks, _ = bw.simulate_walks_array(num_walks=2000, walk_length=3)
    arr = walks.cpu().numpy()
    nbrs = {v: sorted(int(x) for x in wrap.G.neighbors(str(v))) for v in range(n)}
    checked = 0
    for v in range(n):
        if not nbrs[v]:
            continue
        nxt = arr[arr[:, 0] == v, 1]
        counts = np.array([(nxt == x).sum() for x in nbrs[v]])
        checked += _check_freq(counts, np.full(len(nbrs[v]), 1.0 / len(nbrs[v])), f"deepwalk from {v}")
    assert checked >= n // 2
