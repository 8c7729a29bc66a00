"""Host-side text formats (CPU): graph_embs.txt byte-identical to np.savetxt (train.py:193), node2vec .embs.txt
parsed exactly as train.py:79-80 does."""
import numpy as np
import pytest


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_savetxt_is_byte_identical_to_numpy(gdr, tmp_path, dtype):
    from gdr_b200 import textio
    rng = np.random.default_rng(1)
    x = (rng.standard_normal((257, 37)) * 10.0 ** rng.integers(-30, 30, (257, 37))).astype(dtype)
    x[0, :7] = [0.0, -0.0, np.inf, -np.inf, np.nan, np.finfo(dtype).tiny, np.finfo(dtype).max]
    np.savetxt(tmp_path / "ref.txt", x)
    textio.savetxt_e18(tmp_path / "our.txt", x, n_threads=3)
    assert (tmp_path / "ref.txt").read_bytes() == (tmp_path / "our.txt").read_bytes()
    np.savetxt(tmp_path / "ref1.txt", x[:, 0])                 # 1-D input: one value per line
    textio.savetxt_e18(tmp_path / "our1.txt", x[:, 0])
    assert (tmp_path / "ref1.txt").read_bytes() == (tmp_path / "our1.txt").read_bytes()
    textio.savetxt_e18(tmp_path / "empty.txt", np.zeros((0, 5), dtype))
    assert (tmp_path / "empty.txt").read_bytes() == b""


def test_node2vec_embs_parser_matches_python(gdr, tmp_path):
    from gdr_b200 import textio
    rng = np.random.default_rng(2)
    names = [f"DB{i:05d}" if i % 3 else f"node with_{i}".replace(" ", "_") for i in range(501)]
    e = rng.standard_normal((501, 16)).astype(np.float32)
    gdr.synth.write_node2vec_embs(str(tmp_path / "e.txt"), names, e)
    n_ref, v_ref = gdr.synth.read_node2vec_embs(str(tmp_path / "e.txt"))
    n_our, v_our = textio.read_node2vec_embs(str(tmp_path / "e.txt"), n_threads=4)
    assert n_our == n_ref == names
    assert v_our.dtype == np.float64 and np.array_equal(v_our, v_ref)
    (tmp_path / "ragged.txt").write_text("2 3\na 1 2 3\nb 1 2\n")
    with pytest.raises(RuntimeError):
        textio.read_node2vec_embs(str(tmp_path / "ragged.txt"))


def test_weighted_edgelist_is_byte_identical_to_networkx(gdr, tmp_path):
    """predict_drug.py:225-226 writes the MSI with nx.write_weighted_edgelist (weights = repr(float))."""
    import networkx as nx
    from gdr_b200 import textio
    rng = np.random.default_rng(3)
    g = nx.DiGraph()
    names = [f"DB{i:05d}" if i % 2 else str(1000 + i) for i in range(300)] + ["NodeCovid", "GO:0000001"]
    g.add_nodes_from(names)
    specials = [0.1, 1.0, 1e16, 1e15, 123456789012345678.0, 1e-4, 1e-5, 0.00012345, 5e-324, 1.7976931348623157e308,
                3.2071696595616364 / 7, 2.0 ** 53, 1e22, 1.5e-7, 100.0, 0.3, 4.396695660380823 / 3, -2.5, 0.0]
    k = 0
    for _ in range(4000):
        u, v = rng.choice(len(names), 2, replace=False)
        w = specials[k % len(specials)] if k < 3 * len(specials) else float(rng.standard_normal() * 10.0 ** rng.integers(-8, 8))
        g.add_edge(names[u], names[v], weight=w)
        k += 1
    nx.write_weighted_edgelist(g, str(tmp_path / "ref.edgelist"))
    textio.write_weighted_edgelist_graph(g, str(tmp_path / "our.edgelist"), n_threads=3)
    assert (tmp_path / "ref.edgelist").read_bytes() == (tmp_path / "our.edgelist").read_bytes()
    u, v, w = textio.read_weighted_edgelist(str(tmp_path / "ref.edgelist"), n_threads=4)
    h = nx.read_weighted_edgelist(str(tmp_path / "ref.edgelist"), create_using=nx.DiGraph)
    assert len(u) == h.number_of_edges()
    assert all(h[a][b]["weight"] == c for a, b, c in zip(u, v, w))
    (tmp_path / "c.edgelist").write_text("# comment\n\na b 1.5\n  c d 2e-3  \n")
    u, v, w = textio.read_weighted_edgelist(str(tmp_path / "c.edgelist"))
    assert (u, v, w.tolist()) == (["a", "c"], ["b", "d"], [1.5, 0.002])
    (tmp_path / "bad.edgelist").write_text("a b\n")
    with pytest.raises(RuntimeError):
        textio.read_weighted_edgelist(str(tmp_path / "bad.edgelist"))
    textio.write_weighted_edgelist(str(tmp_path / "empty.edgelist"), ["a"], [], [], [])
    assert (tmp_path / "empty.edgelist").read_bytes() == b""


def test_real_msi_edgelist_round_trip(gdr, tmp_path):
    """The real no-PPI MSI (182,204 weighted directed edges): our writer == networkx's, and the reader gets it back."""
    import networkx as nx
    from gdr_b200 import textio
    from .msi_fixture import GoldenMSI
    msi = GoldenMSI("msi_noppi_graph.npz")
    nx.write_weighted_edgelist(msi.graph, str(tmp_path / "ref.edgelist"))
    textio.write_weighted_edgelist_graph(msi.graph, str(tmp_path / "our.edgelist"))
    assert (tmp_path / "ref.edgelist").read_bytes() == (tmp_path / "our.edgelist").read_bytes()
    u, v, w = textio.read_weighted_edgelist(str(tmp_path / "our.edgelist"))
    assert len(u) == 182204
    g = msi.arrays
    names = msi.nodelist
    assert u[:50] == [names[i] for i in g["src"][:50]] or len(set(u)) > 1000
    assert np.array_equal(np.sort(w), np.sort(g["weight"]))
