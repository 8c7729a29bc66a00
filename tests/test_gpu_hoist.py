"""The opt-in reuse of the first layer's aggregations P1 = A_hat X, T1 = A_hat (P1*X) (model.Layer1Cache):
off unless the model is asked, identical results when on, misses after torch-visible writes, explicit invalidation
for writes torch cannot see, and released with the model."""
import gc
import weakref

import numpy as np
import pytest
import scipy.sparse as sp
import torch

pytestmark = pytest.mark.gpu


def _setup(gdr, n=600, d=64, seed=0):
    rng = np.random.default_rng(seed)
    a = sp.random(n, n, density=0.02, random_state=rng, format="csr")
    a = (a + a.T).tocsr()
    adj = gdr.convert_sparse_matrix_to_sparse_tensor(gdr.preprocess_graph(a)).cuda()
    feats = torch.from_numpy(gdr.synth.synth_embeddings(n, d, seed=seed)).cuda()
    np.random.seed(seed)
    torch.manual_seed(seed)
    model = gdr.ResidualGraphConvolutionalNetwork(128, n, num_layers=2, hidden_units=d, layer_decay=0.3).cuda()
    return model, adj, feats


def test_hoist_is_off_by_default_and_sees_data_writes(gdr):
    model, adj, feats = _setup(gdr)
    assert model.hoist_layer1 is False
    h0 = model(x=feats, adj=adj).detach().clone()
    feats.data.copy_(torch.flip(feats, dims=[0]))            # a write torch's version counter does not see
    h1 = model(x=feats, adj=adj).detach()
    assert model._layer1_cache.entry is None and model._layer1_cache.hits == 0
    assert not torch.equal(h0, h1)


def test_hoist_on_is_bit_equal_and_misses_after_inplace_ops(gdr):
    model, adj, feats = _setup(gdr)
    want = model(x=feats, adj=adj).detach().clone()
    model.hoist_layer1 = True
    first = model(x=feats, adj=adj).detach().clone()
    again = model(x=feats, adj=adj).detach().clone()
    assert model._layer1_cache.hits == 1
    assert torch.equal(first, want) and torch.equal(again, want)
    feats.mul_(0.5)                                          # torch-visible write: _version moves, the entry is stale
    got = model(x=feats, adj=adj).detach().clone()
    assert model._layer1_cache.hits == 1
    model.hoist_layer1 = False
    assert torch.equal(got, model(x=feats, adj=adj).detach())


def test_invisible_writes_need_invalidate(gdr):
    model, adj, feats = _setup(gdr)
    model.hoist_layer1 = True
    model(x=feats, adj=adj)
    new = torch.flip(feats, dims=[0]).contiguous()
    feats.data.copy_(new)                                    # not seen by _version ...
    csr = gdr.device_csr_of(adj)
    csr.spmm(new, out=feats)                                 # ... nor is a raw-pointer write by a libgdr kernel
    model.invalidate_layer1_cache()                          # the documented duty of whoever does such writes
    got = model(x=feats, adj=adj).detach().clone()
    assert model._layer1_cache.hits == 0
    model.hoist_layer1 = False
    assert torch.equal(got, model(x=feats, adj=adj).detach())


def test_trainer_hoist_equals_no_hoist_and_cache_dies_with_the_model(gdr):
    from gdr_b200 import trainer as trainer_mod
    results = []
    for hoist in (True, False):
        rng = np.random.default_rng(3)
        n, d = 500, 64
        a = sp.random(n, n, density=0.03, random_state=rng, format="csr")
        adj = gdr.convert_sparse_matrix_to_sparse_tensor(gdr.preprocess_graph((a + a.T).tocsr()))
        feats = gdr.synth.synth_embeddings(n, d, seed=3)
        np.random.seed(3)
        torch.manual_seed(3)
        tr = trainer_mod.GCNTrainer(feats, adj, hidden_units=d, lr=3e-4, batch_size=128, beta=0.3, device="cuda:0",
                                    hoist_layer1=hoist)
        g = torch.Generator().manual_seed(5)
        losses = [tr.iteration(torch.randperm(n, generator=g)[:128].cuda()).item() for _ in range(4)]
        # features handed in explicitly are never cached, whatever the trainer's flag says
        hits = tr.model._layer1_cache.hits
        tr.iteration(torch.arange(128).cuda(), features=tr.features_dev.clone())
        assert tr.model._layer1_cache.hits == hits == (3 if hoist else 0)
        results.append((losses, [p.detach().clone() for p in tr.model.parameters()]))
        if hoist:
            ref = weakref.ref(tr.model._layer1_cache.entry[4])
            csr = gdr.device_csr_of(tr.dataset.adj)
            assert not hasattr(csr, "_layer1_cache")          # nothing pinned on the adjacency any more
            del tr, csr
            gc.collect()
            assert ref() is None
    (l1, p1), (l0, p0) = results
    assert l1 == l0
    assert all(torch.equal(u, v) for u, v in zip(p1, p0))
