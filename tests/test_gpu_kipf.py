"""GPU tests of the Kipf & Welling layer / model mirror (kipf.py over csrc/kipf.cu + spmm.cu + gemm_tc.cu) against the
torch-CPU restatement of multiscale/openne/gcn/{layers,models,metrics}.py (oracle/kipf_oracle.py; the TF1 original cannot
run here, so this half of the parity is "unpinned") and, for the inputs, against the reference-generated utils golden.
Tolerance: north_star's 1e-5 relative or 1e-6 absolute."""
import numpy as np
import pytest
import scipy.sparse as sp
import torch

from oracle import kipf_oracle

from .conftest import assert_close, load_golden

pytestmark = pytest.mark.gpu


def _problem(seed=0, n=300, f=64, classes=7, density=0.03):
    rng = np.random.default_rng(seed)
    a = sp.random(n, n, density=density, random_state=rng, format="csr")
    a = sp.csr_matrix(((a + a.T) > 0).astype(np.float64))
    a.setdiag(0)
    a.eliminate_zeros()
    feats = sp.random(n, f, density=0.1, random_state=rng, format="csr")
    labels = np.zeros((n, classes), dtype=np.float32)
    labels[np.arange(n), rng.integers(0, classes, n)] = 1.0
    mask = np.zeros(n, dtype=bool)
    mask[rng.permutation(n)[: n // 3]] = True
    return a, feats, labels, mask


def _oracle_run(features_t, supports_t, w0, w1, labels, mask, wd, sparse_inputs=True):
    w0o = [w.clone().requires_grad_(True) for w in w0]
    w1o = [w.clone().requires_grad_(True) for w in w1]
    out = kipf_oracle.gcn_forward(features_t, supports_t, w0o, w1o, sparse_inputs=sparse_inputs)
    lb, mk = torch.from_numpy(labels), torch.from_numpy(mask)
    loss = kipf_oracle.gcn_loss(out, lb, mk, w0o, wd)
    loss.backward()
    acc = kipf_oracle.masked_accuracy(out, lb, mk)
    return out.detach(), loss.detach(), acc, [w.grad for w in w0o], [w.grad for w in w1o]


@pytest.mark.parametrize("cheb", [0, 2])
def test_gcn_model_matches_oracle(gdr, cheb):
    """models.py:134-181 end to end: logits, loss, accuracy and all weight gradients; 1 support (preprocess_adj) and 3
    supports (Chebyshev order 2)."""
    kipf = gdr.kipf
    a, feats, labels, mask = _problem()
    ftup = kipf.preprocess_features(feats)
    support = [kipf.preprocess_adj(a)] if cheb == 0 else kipf.chebyshev_polynomials(a, cheb)
    ph = {"support": support, "features": ftup, "labels": labels, "labels_mask": mask, "dropout": 0.0,
          "num_features_nonzero": ftup[1].shape}
    gen = torch.Generator().manual_seed(3)
    model = kipf.GCN(ph, input_dim=feats.shape[1], hidden1=16, weight_decay=5e-4, generator=gen)
    out = model.forward()
    loss = model.loss(out)
    loss.backward()
    acc = model.accuracy(out.detach())
    w0 = [getattr(model.layers[0], f"weights_{i}").detach().cpu() for i in range(len(support))]
    w1 = [getattr(model.layers[1], f"weights_{i}").detach().cpu() for i in range(len(support))]
    o_out, o_loss, o_acc, g0, g1 = _oracle_run(kipf_oracle.to_torch_sparse(ftup),
                                               [kipf_oracle.to_torch_sparse(s) for s in support], w0, w1, labels, mask, 5e-4)
    assert_close(out.detach().cpu().numpy(), o_out.numpy(), what="logits")
    assert_close(loss.item(), o_loss.item(), what="loss")
    assert_close(acc.item(), o_acc.item(), what="masked accuracy")
    for i in range(len(support)):
        assert_close(getattr(model.layers[0], f"weights_{i}").grad.cpu().numpy(), g0[i].numpy(), what=f"dW0_{i}")
        assert_close(getattr(model.layers[1], f"weights_{i}").grad.cpu().numpy(), g1[i].numpy(), what=f"dW1_{i}")
    # predict() = softmax of the outputs (models.py:180-181)
    assert_close(model.predict().detach().cpu().numpy(), torch.softmax(o_out, 1).numpy(), what="predict")


@pytest.mark.parametrize("n,fin,fout,bias,act", [(257, 50, 16, True, "relu"), (400, 128, 128, False, "relu"),
                                                 (130, 33, 7, True, None), (512, 256, 64, True, "tanh")])
def test_graph_convolution_dense_inputs(gdr, n, fin, fout, bias, act):
    """layers.py:165-190 with dense inputs (the second layer's form), input gradient included; odd widths are padded."""
    kipf = gdr.kipf
    a, _, _, _ = _problem(seed=n, n=n)
    support = [kipf.preprocess_adj(a)]
    gen = torch.Generator().manual_seed(n)
    acts = {"relu": torch.relu, None: None, "tanh": torch.tanh}
    layer = kipf.GraphConvolution(fin, fout, {"support": support, "dropout": 0.0}, act=acts[act], bias=bias, generator=gen)
    if bias:
        with torch.no_grad():
            layer.bias.copy_(0.1 * torch.randn(fout, generator=gen))
    x = 0.1 * torch.randn(n, fin, generator=gen)      # magnitudes of normalised features (sums cancel: keep abs errors meaningful)
    xg = x.cuda().requires_grad_(True)
    y = layer(xg)
    up = 0.1 * torch.randn(n, fout, generator=gen)
    (y * up.cuda()).sum().backward()
    xo = x.clone().requires_grad_(True)
    wo = layer.weights_0.detach().cpu().clone().requires_grad_(True)
    bo = layer.bias.detach().cpu().clone().requires_grad_(True) if bias else None
    yo = kipf_oracle.graph_convolution(xo, [kipf_oracle.to_torch_sparse(support[0])], [wo], bias=bo, act=acts[act])
    (yo * up).sum().backward()
    assert_close(y.detach().cpu().numpy(), yo.detach().numpy(), what="output")
    assert_close(layer.weights_0.grad.cpu().numpy(), wo.grad.numpy(), what="dW")
    assert_close(xg.grad.cpu().numpy(), xo.grad.numpy(), what="dX")
    if bias:
        assert_close(layer.bias.grad.cpu().numpy(), bo.grad.numpy(), what="db")


def test_graph_convolution_featureless(gdr):
    kipf = gdr.kipf
    n, fout = 200, 12
    a, _, _, _ = _problem(seed=5, n=n)
    support = [kipf.preprocess_adj(a)]
    layer = kipf.GraphConvolution(n, fout, {"support": support}, featureless=True, act=None,
                                  generator=torch.Generator().manual_seed(1))
    y = layer(None)
    y.sum().backward()
    wo = layer.weights_0.detach().cpu().clone().requires_grad_(True)
    yo = kipf_oracle.graph_convolution(None, [kipf_oracle.to_torch_sparse(support[0])], [wo], act=None, featureless=True)
    yo.sum().backward()
    assert_close(y.detach().cpu().numpy(), yo.detach().numpy(), what="featureless output")
    assert_close(layer.weights_0.grad.cpu().numpy(), wo.grad.numpy(), what="featureless dW")


def test_masked_metrics(gdr):
    """metrics.py:4-21: soft labels, an all-zero-label row, ties in argmax, mask of a single row."""
    kipf = gdr.kipf
    gen = torch.Generator().manual_seed(9)
    n, c = 333, 11
    logits = torch.randn(n, c, generator=gen) * 3
    logits[5, 2] = logits[5, 7] = logits[5].max() + 1.0      # tie: the first maximum wins (tf.argmax)
    labels = torch.rand(n, c, generator=gen)
    labels[:200] = torch.nn.functional.one_hot(torch.randint(0, c, (200,), generator=gen), c).float()
    labels[7] = 0.0
    for mask in (torch.rand(n, generator=gen) < 0.4, torch.arange(n) == 17):
        lg = logits.clone().cuda().requires_grad_(True)
        loss = kipf.masked_softmax_cross_entropy(lg, labels, mask)
        loss.backward()
        lo = logits.clone().requires_grad_(True)
        ref = kipf_oracle.masked_softmax_cross_entropy(lo, labels, mask)
        ref.backward()
        assert_close(loss.item(), ref.item(), what="masked CE")
        assert_close(lg.grad.cpu().numpy(), lo.grad.numpy(), what="dCE/dlogits")
        assert_close(kipf.masked_accuracy(lg.detach(), labels, mask).item(),
                     kipf_oracle.masked_accuracy(logits, labels, mask).item(), what="masked accuracy")


def test_dropout_rule_and_training_step(gdr):
    """TF's dropout rule floor(keep + u)/keep (layers.py:22-29): survivors scaled by 1/keep, expected value preserved;
    sparse_dropout keeps the structure; a few train_step() calls reduce the loss."""
    kipf = gdr.kipf
    x = torch.ones(200, 256, device="cuda", requires_grad=True)
    y = kipf._DropoutFunction.apply(x, 0.8)
    vals = torch.unique(y.detach())
    assert set(np.round(vals.cpu().numpy(), 6).tolist()) <= {0.0, 1.25}
    assert abs(y.mean().item() - 1.0) < 0.02
    y.sum().backward()
    assert torch.equal(x.grad, y.detach())                    # d(x*m)/dx = m, and x = 1
    a, feats, labels, mask = _problem(seed=2)
    ftup = kipf.preprocess_features(feats)
    csr = kipf.as_device_csr(ftup)
    d = kipf.sparse_dropout(csr, 0.5)
    assert d.nnz == csr.nnz and abs((d.val != 0).float().mean().item() - 0.5) < 0.06
    ph = {"support": [kipf.preprocess_adj(a)], "features": ftup, "labels": labels, "labels_mask": mask, "dropout": 0.5}
    torch.manual_seed(0)
    model = kipf.GCN(ph, input_dim=feats.shape[1], hidden1=16, weight_decay=5e-4, generator=torch.Generator().manual_seed(0))
    losses = [model.train_step()[0].item() for _ in range(30)]
    assert np.mean(losses[-5:]) < np.mean(losses[:5])


def test_reference_utils_golden_feeds_the_model(gdr):
    """The arrays the REFERENCE's utils.py produced (tests/golden/kipf_utils.npz) go straight into the model."""
    kipf = gdr.kipf
    g = load_golden("kipf_utils.npz")
    n, f = int(g["n"]), int(g["f"])
    ftup = (g["features_coords"], g["features_values"], tuple(g["features_shape"]))
    support = [(g[f"cheb{i}_coords"], g[f"cheb{i}_values"], tuple(g[f"cheb{i}_shape"])) for i in range(4)]
    rng = np.random.default_rng(0)
    labels = np.zeros((n, 5), dtype=np.float32)
    labels[np.arange(n), rng.integers(0, 5, n)] = 1.0
    ph = {"support": support, "features": ftup, "labels": labels, "labels_mask": g["mask"], "dropout": 0.0}
    model = kipf.GCN(ph, input_dim=f, hidden1=16, weight_decay=5e-4, generator=torch.Generator().manual_seed(4))
    out = model.forward()
    w0 = [getattr(model.layers[0], f"weights_{i}").detach().cpu() for i in range(4)]
    w1 = [getattr(model.layers[1], f"weights_{i}").detach().cpu() for i in range(4)]
    o_out = kipf_oracle.gcn_forward(kipf_oracle.to_torch_sparse(ftup), [kipf_oracle.to_torch_sparse(s) for s in support],
                                    w0, w1)
    assert_close(out.detach().cpu().numpy(), o_out.numpy(), what="logits on the reference's supports")
