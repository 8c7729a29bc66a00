"""CPU tests of the drop-in boundary: libgdr.so loads, exports every symbol include/gdr.h
declares, argument errors come back as negative codes with a message, and the product package
refuses to run without a CUDA device (no CPU fallback).  No compute call is made here.
"""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gdr.h")


def header_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gdr_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_all_exported(gdr):
    lib = ctypes.CDLL(gdr._ffi.LIB_PATH)
    names = header_symbols()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/gdr.h but not exported by libgdr.so: {missing}"


def test_ffi_table_covers_the_header(gdr):
    assert sorted(gdr._ffi.EXPORTED_SYMBOLS) == header_symbols()


def test_version_and_launch_counter(gdr):
    L = gdr._ffi.lib()
    assert L.gdr_version() >= 100
    L.gdr_launch_count_reset()
    assert L.gdr_launch_count() == 0


def test_argument_errors_do_not_launch(gdr):
    L = gdr._ffi.lib()
    rc = L.gdr_spmm_csr_f32(None, None, None, 10, 128, None, None, None, None, None, None)
    assert rc == -1 and b"null" in L.gdr_last_error_string()
    rc = L.gdr_spmm_csr_f32(None, None, None, 10, 126, None, None, None, None, None, None)
    assert rc == -2
    assert L.gdr_csr_from_coo_workspace_bytes(-1, 1, 1) < 0
    assert L.gdr_csr_from_coo_workspace_bytes(1000, 10, 10) > 0
    assert L.gdr_launch_count() == 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback(gdr):
    import scipy.sparse as sp
    with pytest.raises(RuntimeError, match="CUDA"):
        gdr.preprocess_graph(sp.identity(4, format="csr"))
    model = gdr.ResidualGraphConvolutionalNetwork(1, 1, num_layers=2, hidden_units=8)
    adj = gdr.convert_sparse_matrix_to_sparse_tensor(sp.identity(4, format="csr"))
    with pytest.raises(RuntimeError):
        model(torch.randn(4, 8), adj)
    with pytest.raises(RuntimeError):
        gdr.GSS_loss(1.0).gss_loss(torch.randn(4, 8), 0.1)


def test_dataset_and_loader_follow_reference(gdr):
    feats = torch.arange(40, dtype=torch.float64).reshape(10, 4)
    ds = gdr.DiffusionDataSet(feats, adj="adj")
    assert ds.features.dtype == torch.float32 and len(ds) == 10 and ds[3] == 3 and ds.adj == "adj"
    torch.manual_seed(0)
    got = [b.tolist() for b in gdr.DiffusionDataLoader(ds, batch_size=4, num_workers=6, shuffle=True)]
    torch.manual_seed(0)
    exp = [b.tolist() for b in torch.utils.data.DataLoader(ds, batch_size=4, shuffle=True)]
    assert got == exp and [len(b) for b in got] == [4, 4, 2]
