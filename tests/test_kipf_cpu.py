"""CPU tests of the Kipf-GCN utils mirror (gcn-drug-repurposing_b200/kipf.py) and of its oracle against the golden
arrays that oracle/gen_golden_kipf.py produced from the REFERENCE's multiscale/openne/gcn/utils.py."""
import numpy as np
import scipy.sparse as sp

from oracle import kipf_oracle

from .conftest import load_golden


def _inputs(g):
    n, f = int(g["n"]), int(g["f"])
    a = sp.csr_matrix((g["adj_data"], g["adj_indices"], g["adj_indptr"]), shape=(n, n))
    feats = sp.coo_matrix((g["feat_data"], (g["feat_row"], g["feat_col"])), shape=(n, f))
    return a, feats


def _same(tup, g, name, exact=True):
    coords, values, shape = tup
    assert np.array_equal(np.asarray(coords), g[f"{name}_coords"])
    assert tuple(shape) == tuple(g[f"{name}_shape"])
    if exact:
        assert np.array_equal(np.asarray(values), g[f"{name}_values"])
    else:
        np.testing.assert_allclose(np.asarray(values), g[f"{name}_values"], rtol=1e-12, atol=1e-14)


def test_oracle_utils_match_reference_golden():
    g = load_golden("kipf_utils.npz")
    a, feats = _inputs(g)
    _same(kipf_oracle.preprocess_features(sp.lil_matrix(feats)), g, "features")
    _same(kipf_oracle.preprocess_adj(a), g, "support")
    for i, t in enumerate(kipf_oracle.chebyshev_polynomials(a, 3)):
        _same(t, g, f"cheb{i}", exact=False)          # eigsh starts from a random vector: last-bit differences upstream too


def test_product_utils_match_reference_golden():
    import importlib
    kipf = importlib.import_module("gcn-drug-repurposing_b200.kipf")
    g = load_golden("kipf_utils.npz")
    a, feats = _inputs(g)
    _same(kipf.preprocess_features(sp.lil_matrix(feats)), g, "features")
    _same(kipf.preprocess_adj(a), g, "support")
    for i, t in enumerate(kipf.chebyshev_polynomials(a, 3)):
        _same(t, g, f"cheb{i}", exact=False)
    assert np.array_equal(kipf.sample_mask(range(10, 30), int(g["n"])), g["mask"])
