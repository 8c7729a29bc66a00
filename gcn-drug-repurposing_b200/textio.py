"""Text formats on either side of the path, formatted / parsed by libgdr's host code:

    savetxt_e18(path, x)      byte-identical to np.savetxt(path, x) (train.py:193: '%.18e', ' ', '\\n', no header)
    read_node2vec_embs(path)  the (names, fp64 [N, d]) that train.py:79-80 extracts from a node2vec .embs.txt

Both are host I/O, not compute: they exist because after the kernels np.savetxt / per-token Python parsing are the
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
