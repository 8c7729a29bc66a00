"""GPU tests of the tensor-core kNN selection (csrc/gram_topk_tc.cu, gdr_gram_topk_tc_f64) through the C ABI.

What is checked: the tcgen05 path (TF32 Gram tiles + fused running top-C + exact fp64 re-rank + certified fallback)
returns the SAME index sets and fp64 values as the exact fp64 kernel (gdr_gram_topk_f64), as numpy's fp64
`np.matmul` + `argpartition` (helpers/helper.py:39,44), and as the reference-generated goldens; north_star:
"kNN neighbour index sets must be bit-exact, ties excepted"."""
import numpy as np
import pytest
import torch

from .conftest import load_golden

pytestmark = pytest.mark.gpu


def _both(gdr, x, k):
    xt = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    i_tc, v_tc = gdr.helper.gram_topk(xt, k, backend="tc")
    nfb = int(gdr.helper.last_gram_topk_info["fallback_rows"].item())
    i_64, v_64 = gdr.helper.gram_topk(xt, k, backend="f64")
    return i_tc.cpu().numpy(), v_tc.cpu().numpy(), i_64.cpu().numpy(), v_64.cpu().numpy(), nfb


@pytest.mark.parametrize("n,d,k", [(64, 8, 5), (65, 20, 1), (129, 32, 3), (1000, 128, 5), (300, 130, 5), (5000, 100, 8),
                                   (3000, 256, 16), (2049, 64, 9)])
def test_gram_topk_tc_equals_fp64_kernel_and_numpy(gdr, n, d, k):
    x = np.random.default_rng(n + d).standard_normal((n, d))
    x *= np.random.default_rng(n).uniform(0.5, 3.0, size=(n, 1))      # raw node2vec rows are not unit-norm
    i_tc, v_tc, i_64, v_64, nfb = _both(gdr, x, k)
    assert np.array_equal(v_tc, v_64), "fp64 values differ from the exact kernel"
    assert np.array_equal(i_tc, i_64), "index lists differ from the exact kernel"
    s = x @ x.T
    exp_val = -np.sort(-s, axis=1)[:, :k]
    np.testing.assert_allclose(v_tc, exp_val, rtol=1e-12, atol=1e-12)
    exp_set = np.sort(np.argpartition(s, -k, 1)[:, -k:], axis=1)      # helpers/helper.py:44
    gap = exp_val[:, -1] - (-np.sort(-s, axis=1)[:, k] if n > k else -np.inf)
    clear = gap > 1e-12
    assert np.array_equal(np.sort(i_tc, axis=1)[clear], exp_set[clear])
    assert nfb <= max(1, n // 50), f"{nfb} of {n} rows took the fp64 path on well-separated data"


@pytest.mark.parametrize("case", ["gcn_knn257.npz", "gcn_knn600_l3.npz", "gcn_knn2000.npz"])
def test_gram_topk_tc_matches_reference_goldens(gdr, case):
    g = load_golden(case)
    k = int(g["k"])
    x = g["embs"].astype(np.float64)
    i_tc, v_tc, i_64, v_64, _ = _both(gdr, x, k)
    clear = g["topk_gap"] > 1e-12
    assert np.array_equal(np.sort(i_tc, axis=1)[clear], g["topk_sorted"][clear])
    assert np.array_equal(i_tc, i_64) and np.array_equal(v_tc, v_64)


def test_gram_topk_tc_near_ties_take_the_exact_path(gdr):
    """Clusters of near-identical rows: more near-ties inside the TF32 margin than a candidate list holds, so rows
    cannot be certified and must come out of the fp64 fallback — with the same answer."""
    rng = np.random.default_rng(7)
    centres = rng.standard_normal((4, 64))
    x = np.repeat(centres, 600, axis=0) + 1e-7 * rng.standard_normal((2400, 64))
    x = x[rng.permutation(len(x))]
    i_tc, v_tc, i_64, v_64, nfb = _both(gdr, x, 5)
    assert nfb > 0, "expected uncertified rows"
    assert np.array_equal(v_tc, v_64)
    s = x @ x.T
    np.testing.assert_allclose(v_tc, -np.sort(-s, axis=1)[:, :5], rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(np.take_along_axis(s, i_tc.astype(np.int64), 1), v_tc, rtol=1e-12, atol=1e-12)


def test_gram_topk_tc_duplicate_rows_and_zero_rows(gdr):
    x = np.random.default_rng(3).standard_normal((500, 128))
    x[100:110] = x[0:10]          # exact duplicates: genuine fp64 ties
    x[200:205] = 0.0              # zero rows: every similarity is 0
    i_tc, v_tc, i_64, v_64, _ = _both(gdr, x, 5)
    assert np.array_equal(v_tc, v_64)
    s = x @ x.T
    np.testing.assert_allclose(np.take_along_axis(s, i_tc.astype(np.int64), 1), v_tc, rtol=1e-12, atol=1e-12)
    assert all(len(set(r)) == 5 for r in i_tc.tolist())


@pytest.mark.parametrize("factor", [1e-25, 1e25, 2.0 ** -300, 2.0 ** 300])
def test_gram_topk_tc_extreme_scales(gdr, factor):
    """Entries far outside the fp32 range (1e25^2 overflows fp32, 1e-25^2 underflows it): the kernel rescales by a
    power of two before the TF32 product, so the index sets still equal the fp64 kernel's and rows stay certified."""
    x = np.random.default_rng(11).standard_normal((1500, 96)) * factor
    i_tc, v_tc, i_64, v_64, nfb = _both(gdr, x, 5)
    assert np.isfinite(v_tc).all()
    assert np.array_equal(v_tc, v_64)
    assert np.array_equal(np.sort(i_tc, axis=1), np.sort(i_64, axis=1))
    assert nfb < 150, f"{nfb} rows fell back: the rescaling is not effective"


def test_gram_topk_tc_rows_of_wildly_different_scale(gdr):
    """Some rows 1e-30 times smaller than the rest: their TF32 operands fall below the fp32 normal range, the
    relative error bound no longer covers them, and the absolute term of the margin sends them to the fp64 path."""
    x = np.random.default_rng(12).standard_normal((1200, 64))
    x[::7] *= 1e-42
    i_tc, v_tc, i_64, v_64, nfb = _both(gdr, x, 5)
    assert np.array_equal(v_tc, v_64)
    s = x @ x.T
    np.testing.assert_allclose(np.take_along_axis(s, i_tc.astype(np.int64), 1), v_tc, rtol=1e-12, atol=0)


def test_gram_topk_tc_msi_scale(gdr):
    """N = 29,960, d = 128, k = 5 (SURVEY 8 'MSI-kNN'): full equality with the exact kernel."""
    n, d, k = 29960, 128, 5
    x = gdr.synth.synth_embeddings(n, d, seed=5).astype(np.float64)
    i_tc, v_tc, i_64, v_64, nfb = _both(gdr, x, k)
    assert np.array_equal(v_tc, v_64)
    same = (i_tc == i_64).all(axis=1)
    assert same.mean() > 0.9999       # exact fp64 ties may order differently
    # the rows that differ: every index either kernel returned must really carry the value returned with it (so the
    # two answers differ only by a permutation inside, or a swap at the boundary of, a genuine fp64 tie)
    for r in np.flatnonzero(~same):
        for idx, val in ((i_tc[r], v_tc[r]), (i_64[r], v_64[r])):
            assert len(set(idx.tolist())) == k
            np.testing.assert_allclose(x[idx.astype(np.int64)] @ x[r], val, rtol=1e-12, atol=0)
    print(f"msi-scale gram top-k: {nfb} of {n} rows took the fp64 path")
    assert nfb < n // 100


def test_gram_topk_tc_argument_errors(gdr):
    x = torch.zeros(10, 512, dtype=torch.float64, device="cuda")
    with pytest.raises(RuntimeError, match="exceeds 256"):
        gdr.helper.gram_topk(x, 5, backend="tc")
    with pytest.raises(RuntimeError, match="outside 1..16"):
        gdr.helper.gram_topk(x[:, :64].contiguous(), 17, backend="tc")
    with pytest.raises(RuntimeError, match="exceeds n"):
        gdr.helper.gram_topk(x[:3, :64].contiguous(), 5, backend="tc")
