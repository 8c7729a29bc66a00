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
