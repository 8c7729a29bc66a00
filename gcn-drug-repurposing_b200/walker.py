"""Drop-in mirror of the walk generators of the reference's vendored OpenNE — SURVEY.md 8(f) row 4:

    Walker(G, p, q, workers)          multiscale/openne/walker.py:58-161   node2vec second-order walks
    BasicWalker(G, workers)           multiscale/openne/walker.py:11-55    DeepWalk (uniform neighbour choice)

`G` is the reference's graph wrapper (multiscale/openne/graph.py): `.G` a networkx DiGraph with 'weight' on every edge,
`.node_size`, `.look_up_dict`.  Same methods: `preprocess_transition_probs()` then `simulate_walks(num_walks,
walk_length)` -> list of walks, each a list of node ids, nodes re-shuffled with Python's `random` before every pass
exactly as the reference does (so `random.seed(s)` fixes the start order of both).  The steps themselves run on the
GPU (csrc/walk.cu): first-order draws by inverse CDF over the row's running weight sums, second-order steps by
rejection against max(1/p, 1, 1/q) — the same transition probabilities as the reference's per-edge alias tables
without their O(sum deg^2) memory.  The random stream is the device's counter-based generator, not numpy's: parity with
the reference is distributional (tests/test_gpu_walk.py), as SURVEY 8(f) states.
"""
import random

import numpy as np
import torch

from . import _ffi


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("gcn-drug-repurposing_b200 needs a CUDA device: there is no CPU implementation")
    return torch.device("cuda", torch.cuda.current_device())


class _DeviceWalkGraph:
    """CSR of the DiGraph in list(G.nodes()) order, columns ascending inside a row, fp32 weights + running sums."""

    def __init__(self, nxg, weighted):
        self.nodes = list(nxg.nodes())
        index = {v: i for i, v in enumerate(self.nodes)}
        n = len(self.nodes)
        src, dst, wts = [], [], []
        for u, v, w in nxg.edges(data="weight", default=1.0):
            src.append(index[u])
            dst.append(index[v])
            wts.append(float(w))
        src = np.asarray(src, dtype=np.int64)
        dst = np.asarray(dst, dtype=np.int64)
        wts = np.asarray(wts, dtype=np.float32)
        order = np.lexsort((dst, src))
        src, dst, wts = src[order], dst[order], wts[order]
        rowptr = np.zeros(n + 1, dtype=np.int64)
        np.add.at(rowptr, src + 1, 1)
        rowptr = np.cumsum(rowptr)
        dev = _device()
        self.n = n
        self.rowptr = torch.from_numpy(rowptr.astype(np.int32)).to(dev)
        self.col = torch.from_numpy(dst.astype(np.int32)).to(dev)
        self.weight = torch.from_numpy(wts).to(dev) if weighted else None
        self.cum = torch.empty(max(len(dst), 1), dtype=torch.float32, device=dev)
        _ffi.check(_ffi.lib().gdr_walk_row_cumsum_f32(_ffi.ptr(self.rowptr), _ffi.ptr(self.weight), n, _ffi.ptr(self.cum),
                                                      _ffi.stream()), "walk_row_cumsum")


class _WalkerBase:
    second_order = False
    weighted = True

    def _setup(self, G):
        self.G = G.G
        self.node_size = getattr(G, "node_size", len(G.G))
        self.look_up_dict = getattr(G, "look_up_dict", None)
        self._dev_graph = None

    def _graph(self):
        if self._dev_graph is None:
            self._dev_graph = _DeviceWalkGraph(self.G, self.weighted)
        return self._dev_graph

    def simulate_walks_array(self, num_walks, walk_length):
        """All walks as one int32 array [num_walks * N, walk_length] of node INDICES (position in list(G.nodes())),
        -1 after a walk has ended at a node without out-neighbours; plus the node list."""
        g = self._graph()
        L = _ffi.lib()
        order = list(range(g.n))
        walks = torch.empty(num_walks * g.n, walk_length, dtype=torch.int32, device=g.rowptr.device)
        # walker.py:47,102: a new node order every pass.  The reference's node2vec walker takes nothing else from
        # Python's `random` (alias_draw uses np.random), so all shuffles are drawn first, back to back, and only then
        # the seed of the device stream: after random.seed(s) every pass starts in the reference's order.
        orders = []
        for _ in range(num_walks):
            random.shuffle(order)
            orders.append(list(order))
        seed = random.getrandbits(63)
        print('Walk iteration:')
        for it in range(num_walks):
            print(str(it + 1), '/', str(num_walks))
            start = torch.tensor(orders[it], dtype=torch.int32, device=walks.device)
            _ffi.check(L.gdr_node2vec_walks(_ffi.ptr(g.rowptr), _ffi.ptr(g.col), _ffi.ptr(g.cum), g.n, _ffi.ptr(start),
                                            g.n, walk_length, float(getattr(self, "p", 1.0)), float(getattr(self, "q", 1.0)),
                                            1 if self.second_order else 0, seed, it * g.n,
                                            _ffi.ptr(walks[it * g.n:(it + 1) * g.n]), _ffi.stream()), "node2vec_walks")
        return walks, g.nodes

    def simulate_walks(self, num_walks, walk_length):
        """walker.py:36-55 / 93-106: list of walks (lists of node ids)."""
        walks, nodes = self.simulate_walks_array(num_walks, walk_length)
        arr = walks.cpu().numpy()
        names = np.asarray(nodes, dtype=object)
        lengths = np.where((arr < 0).any(1), (arr < 0).argmax(1), arr.shape[1])
        return [names[row[:ln]].tolist() for row, ln in zip(arr, lengths)]


class BasicWalker(_WalkerBase):
    """walker.py:11-55: `random.choice(cur_nbrs)` — uniform over the out-neighbours, edge weights ignored."""
    second_order = False
    weighted = False

    def __init__(self, G, workers):
        self._setup(G)

    def deepwalk_walk(self, walk_length, start_node):
        g = self._graph()
        return self._one(g, walk_length, start_node)

    def _one(self, g, walk_length, start_node):
        L = _ffi.lib()
        start = torch.tensor([g.nodes.index(start_node)], dtype=torch.int32, device=g.rowptr.device)
        out = torch.empty(1, walk_length, dtype=torch.int32, device=start.device)
        _ffi.check(L.gdr_node2vec_walks(_ffi.ptr(g.rowptr), _ffi.ptr(g.col), _ffi.ptr(g.cum), g.n, _ffi.ptr(start), 1,
                                        walk_length, float(getattr(self, "p", 1.0)), float(getattr(self, "q", 1.0)),
                                        1 if self.second_order else 0, random.getrandbits(63), 0, _ffi.ptr(out),
                                        _ffi.stream()), "node2vec_walks")
        row = out[0].cpu().numpy()
        return [g.nodes[i] for i in row[row >= 0]]


class Walker(BasicWalker):
    """walker.py:58-161: node2vec with return parameter p and in-out parameter q."""
    second_order = True
    weighted = True

    def __init__(self, G, p, q, workers):
        self._setup(G)
        self.p, self.q = p, q

    def preprocess_transition_probs(self):
        """walker.py:127-161 builds alias_nodes / alias_edges; here only the first-order tables exist (running weight sums
        per row, built on the device) — the second-order distribution is sampled by rejection inside the walk."""
        self._graph()

    def node2vec_walk(self, walk_length, start_node):
        return self._one(self._graph(), walk_length, start_node)
