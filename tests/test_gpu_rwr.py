"""GPU parity tests of the diffusion-profile path (csrc/rwr.cu through gdr_rwr_solve_f64) against the
golden profiles produced by the reference's DiffusionProfiles (oracle/gen_golden.py) and against the
oracle restatement (oracle/rwr_oracle.py).  fp64: profiles are compared at 1e-12 relative (north_star
asks 1e-5), iteration counts exactly."""
import os

import numpy as np
import pytest

from .conftest import load_golden
from oracle import rwr_oracle

pytestmark = pytest.mark.gpu


class SynthMSI:
    """Duck-typed stand-in for multiscale/msi/msi.py's MSI: exactly the attributes DiffusionProfiles reads."""

    class _Graph:
        def __init__(self, g):
            self._g = g

        def edges(self, data=None, default=None):
            names = self._g["names"]
            return ((names[u], names[v], float(w)) for u, v, w in zip(self._g["src"], self._g["dst"], self._g["weight"]))

        def __len__(self):
            return self._g["n"]

    def __init__(self, g, synth):
        names = g["names"]
        self.graph = self._Graph(g)
        self.nodelist = list(names)
        self.node2idx = {nm: i for i, nm in enumerate(names)}
        t = g["node_type"]
        self.drugs_in_graph = [names[i] for i in g["seeds"] if t[i] == synth.DRUG]
        self.indications_in_graph = [names[i] for i in g["seeds"] if t[i] == synth.INDICATION]
        self.drug_or_indication2proteins = {
            names[s]: set(names[p] for p in g["seed_proteins"].get(int(s), [])) for s in g["seeds"]}
        self.saved = []

    def save_graph(self, path):
        self.saved.append("graph")

    def save_node2idx(self, path):
        self.saved.append("node2idx")

    def weight_graph(self, weights):
        self.saved.append("weighted")


def _setup(gdr, scale, seed):
    g = gdr.synth.msi_synth(seed=seed, scale=scale)
    msi = SynthMSI(g, gdr.synth)
    dp = gdr.DiffusionProfiles(alpha=gdr.synth.RWR_ALPHA, max_iter=gdr.synth.RWR_MAX_ITER, tol=gdr.synth.RWR_TOL,
                               weights=None, num_cores=4, save_load_file_path=None)
    dp.get_initial_M(msi)
    return g, msi, dp


def test_rwr_matches_reference_golden(gdr):
    gold = load_golden("rwr_msi600.npz")
    g, msi, dp = _setup(gdr, float(gold["scale"]), int(gold["seed"]))
    m0 = rwr_oracle.initial_matrix(g["n"], g["src"], g["dst"], g["weight"])
    assert (dp.initial_M != m0).nnz == 0
    names = [g["names"][int(g["seeds"][c])] for c in gold["chosen"]]
    x, iters = dp.solve(msi, names)
    assert np.array_equal(iters, gold["iters"])                      # same stopping iteration per seed
    np.testing.assert_allclose(x, gold["profiles"], rtol=1e-12, atol=1e-300)
    assert np.abs(x.sum(axis=1) - 1.0).max() < 1e-9
    # bit-for-bit on (nearly) every entry: the accumulation order of the reference is reproduced
    assert (x == gold["profiles"]).mean() > 0.999


def test_rwr_all_seeds_vs_oracle_and_files(gdr, tmp_path):
    g, msi, dp = _setup(gdr, 0.01, 11)
    dp.save_load_file_path = str(tmp_path)
    dp.calculate_diffusion_profiles(msi)
    assert msi.saved == ["graph", "node2idx", "weighted"]
    seeds = msi.drugs_in_graph + msi.indications_in_graph
    dp.load_diffusion_profiles(seeds)
    assert set(dp.drug_or_indication2diffusion_profile) == set(seeds)
    m0 = rwr_oracle.initial_matrix(g["n"], g["src"], g["dst"], g["weight"])
    sinks = [int(s) for s in g["seeds"]]
    for nm in seeds[::3]:
        ox, _ = rwr_oracle.diffusion_profile(m0, msi.node2idx[nm], g["seed_proteins"], sinks, gdr.synth.RWR_ALPHA,
                                             gdr.synth.RWR_TOL, gdr.synth.RWR_MAX_ITER)
        got = dp.drug_or_indication2diffusion_profile[nm]
        assert got.dtype == np.float64 and got.shape == (g["n"],)
        np.testing.assert_allclose(got, ox, rtol=1e-12, atol=1e-300)
    assert os.path.exists(os.path.join(str(tmp_path), dp.clean_file_name(seeds[0]) + "_p_visit_array.npy"))


def test_rwr_max_iter_raises(gdr):
    g, msi, dp = _setup(gdr, 0.01, 3)
    dp.max_iter = 2
    with pytest.raises(RuntimeError, match="failed to converge"):
        dp.solve(msi, msi.drugs_in_graph[:5])
