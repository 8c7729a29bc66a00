"""Generates tests/golden/kipf_utils.npz from the REFERENCE's own multiscale/openne/gcn/utils.py (imported from
/root/reference in the build container; the module needs one shim: scipy.sparse.linalg.eigen.arpack moved in
scipy >= 1.8) and asserts that oracle/kipf_oracle.py reproduces every array bit for bit.

    python oracle/gen_golden_kipf.py
"""
import importlib.util
import os
import sys
import types

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import kipf_oracle  # noqa: E402

REF = "/root/reference/multiscale/openne/gcn/utils.py"


def load_reference_utils():
    eigen = types.ModuleType("scipy.sparse.linalg.eigen")
    arpack = types.ModuleType("scipy.sparse.linalg.eigen.arpack")
    arpack.eigsh = scipy.sparse.linalg.eigsh
    eigen.arpack = arpack
    sys.modules["scipy.sparse.linalg.eigen"] = eigen
    sys.modules["scipy.sparse.linalg.eigen.arpack"] = arpack
    if not hasattr(np, "bool"):
        np.bool = bool
    spec = importlib.util.spec_from_file_location("ref_gcn_utils", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ref = load_reference_utils()
    rng = np.random.default_rng(11)
    n, f = 90, 40
    a = sp.random(n, n, density=0.05, random_state=rng, format="csr")
    a = ((a + a.T) > 0).astype(np.float64)
    a.setdiag(0)
    a.eliminate_zeros()
    a = sp.csr_matrix(a)
    feats = sp.random(n, f, density=0.1, random_state=rng, format="lil")
    feats[5, :] = 0                                       # an empty feature row (r_inv = inf -> 0)
    out = {"adj_indptr": a.indptr, "adj_indices": a.indices, "adj_data": a.data, "n": n, "f": f}
    fc = sp.coo_matrix(feats)
    out.update(feat_row=fc.row, feat_col=fc.col, feat_data=fc.data)

    def put(name, tup, otup, exact=True):
        for part, x, y in zip(("coords", "values", "shape"), tup, otup):
            x, y = np.asarray(x), np.asarray(y)
            assert x.shape == y.shape, f"oracle != reference for {name}.{part} (shape)"
            if exact or part != "values":
                assert np.array_equal(x, y), f"oracle != reference for {name}.{part}"
            else:
                # eigsh starts ARPACK from a random vector: the reference itself differs from run to run in the
                # last bits of lambda_max (measured: 4e-16 .. 9e-16 on these values), so values are pinned to 1e-12
                assert np.allclose(x, y, rtol=1e-12, atol=1e-14), f"oracle != reference for {name}.{part}"
            out[f"{name}_{part}"] = x

    put("features", ref.preprocess_features(sp.lil_matrix(feats)), kipf_oracle.preprocess_features(sp.lil_matrix(feats)))
    put("support", ref.preprocess_adj(a), kipf_oracle.preprocess_adj(a))
    rc = ref.chebyshev_polynomials(a, 3)
    oc = kipf_oracle.chebyshev_polynomials(a, 3)
    assert len(rc) == len(oc) == 4
    for i, (r, o) in enumerate(zip(rc, oc)):
        put(f"cheb{i}", r, o, exact=False)
    m = ref.sample_mask(range(10, 30), n)
    out["mask"] = np.asarray(m)
    path = os.path.join(ROOT, "tests", "golden", "kipf_utils.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    main()
