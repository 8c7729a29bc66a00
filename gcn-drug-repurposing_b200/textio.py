"""Text formats on either side of the path, formatted / parsed by libgdr's host code:

    savetxt_e18(path, x)      byte-identical to np.savetxt(path, x) (train.py:193: '%.18e', ' ', '\\n', no header)
    read_node2vec_embs(path)  the (names, fp64 [N, d]) that train.py:79-80 extracts from a node2vec .embs.txt

    write_weighted_edgelist   byte-identical to nx.write_weighted_edgelist(G, path) (predict_drug.py:225-226)
    read_weighted_edgelist    the (u names, v names, fp64 weights) nx.read_weighted_edgelist would load (run_covid.py:150)

All are host I/O, not compute: they exist because after the kernels np.savetxt / per-token Python parsing are the
longest steps of a train.py run at MSI scale."""
import ctypes

import numpy as np

from . import _ffi


def savetxt_e18(path, x, n_threads=0):
    x = np.ascontiguousarray(np.asarray(x))
    if x.ndim == 1:
        x = x.reshape(-1, 1)                                   # np.savetxt writes a 1-D array as one column
    if x.ndim != 2 or x.dtype not in (np.float32, np.float64):
        raise RuntimeError("savetxt_e18: a 2-D float32 / float64 array is expected")
    L = _ffi.lib()
    n, d = x.shape
    cap = int(L.gdr_text_savetxt_bytes(n, d))
    buf = ctypes.create_string_buffer(cap)
    w = L.gdr_text_savetxt_e18(x.ctypes.data, n, d, x.dtype.itemsize, buf, cap, n_threads)
    if w < 0:
        _ffi.check(int(w), "text_savetxt_e18")
    with open(path, "wb") as f:
        f.write(memoryview(buf)[:w])


def read_node2vec_embs(path, n_threads=0):
    L = _ffi.lib()
    with open(path, "rb") as f:
        text = f.read()
    d = ctypes.c_int64(0)
    n = L.gdr_text_parse_embs(text, len(text), 1, None, 0, None, None, ctypes.byref(d), n_threads)
    if n < 0:
        _ffi.check(int(n), "text_parse_embs")
    vals = np.empty((n, d.value), dtype=np.float64)
    off = np.empty(n, dtype=np.int64)
    ln = np.empty(n, dtype=np.int32)
    rc = L.gdr_text_parse_embs(text, len(text), 1, vals.ctypes.data, d.value, off.ctypes.data, ln.ctypes.data, None,
                               n_threads)
    if rc < 0:
        _ffi.check(int(rc), "text_parse_embs")
    names = [text[o:o + k].decode() for o, k in zip(off.tolist(), ln.tolist())]
    return names, vals


def write_weighted_edgelist(path, names, src, dst, weight, n_threads=0):
    """names: node names (str) by node id; src/dst: int arrays of node ids in the order the edges are to be written
    (nx.write_weighted_edgelist writes G.edges(data=True) order); weight: fp64.  One line 'u v repr(w)' per edge."""
    L = _ffi.lib()
    enc = [str(x).encode("utf-8") for x in names]
    ln = np.array([len(b) for b in enc], dtype=np.int32)
    off = np.zeros(len(enc), dtype=np.int64)
    if len(enc) > 1:
        np.cumsum(ln[:-1], out=off[1:])
    blob = b"".join(enc)
    src = np.ascontiguousarray(src, dtype=np.int32)
    dst = np.ascontiguousarray(dst, dtype=np.int32)
    w = np.ascontiguousarray(weight, dtype=np.float64)
    args = (blob, off.ctypes.data, ln.ctypes.data, len(enc), src.ctypes.data, dst.ctypes.data, w.ctypes.data, len(src))
    need = L.gdr_text_format_edgelist(*args, None, 0, n_threads)
    if need > 0 or (need < 0 and need > -6 and len(src) == 0):
        _ffi.check(int(need), "text_format_edgelist")
    cap = int(-need) if need < 0 else 0
    buf = ctypes.create_string_buffer(max(cap, 1))
    wrote = L.gdr_text_format_edgelist(*args, buf, cap, n_threads) if cap else 0
    if wrote < 0:
        _ffi.check(int(wrote), "text_format_edgelist")
    with open(path, "wb") as f:
        f.write(memoryview(buf)[:wrote])


def write_weighted_edgelist_graph(G, path, n_threads=0):
    """nx.write_weighted_edgelist(G, path) for a networkx (Di)Graph whose edges carry a float 'weight'."""
    names = list(G.nodes())
    idx = {n: i for i, n in enumerate(names)}
    e = [(idx[u], idx[v], w) for u, v, w in G.edges(data="weight")]
    a = np.array(e, dtype=np.float64).reshape(-1, 3)
    write_weighted_edgelist(path, names, a[:, 0].astype(np.int32), a[:, 1].astype(np.int32), a[:, 2], n_threads)


def read_weighted_edgelist(path, n_threads=0):
    """(u names, v names, fp64 weights), one entry per non-comment line, in file order."""
    L = _ffi.lib()
    with open(path, "rb") as f:
        text = f.read()
    n = L.gdr_text_parse_edgelist(text, len(text), None, None, None, None, None, n_threads)
    if n < 0:
        _ffi.check(int(n), "text_parse_edgelist")
    uo, vo = np.empty(n, dtype=np.int64), np.empty(n, dtype=np.int64)
    ul, vl = np.empty(n, dtype=np.int32), np.empty(n, dtype=np.int32)
    w = np.empty(n, dtype=np.float64)
    rc = L.gdr_text_parse_edgelist(text, len(text), uo.ctypes.data, ul.ctypes.data, vo.ctypes.data, vl.ctypes.data,
                                   w.ctypes.data, n_threads)
    if rc < 0:
        _ffi.check(int(rc), "text_parse_edgelist")
    u = [text[o:o + k].decode() for o, k in zip(uo.tolist(), ul.tolist())]
    v = [text[o:o + k].decode() for o, k in zip(vo.tolist(), vl.tolist())]
    return u, v, w
