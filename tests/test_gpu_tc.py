"""GPU tests of the tcgen05 (3xTF32) GEMM family (csrc/gemm_tc.cu) through the C ABI.

Oracle: the same contraction in fp64 on the host (numpy), and the exact-fp32 SIMT kernels of this library /
torch CPU fp32 as the reference-arithmetic yardstick.  Tolerance (north_star): <= 1e-5 relative or <= 1e-6
absolute on fp32 activations / gradients; in addition the 3xTF32 error against fp64 must stay within 4x the
error the reference's own fp32 arithmetic has against fp64 (i.e. it is fp32-grade, not TF32-grade)."""
import numpy as np
import pytest
import torch

from .conftest import assert_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _gemm_nt_tc(gdr, a, b):
    L = gdr._ffi.lib()
    m, k = a.shape
    n = b.shape[0]
    c = torch.full((m, n), float("nan"), device=DEV)
    ws = gdr._ffi.workspace(L.gdr_gemm_nt_tc_workspace_bytes(m, n, k), a.device)
    gdr._ffi.check(L.gdr_gemm_nt_tc_f32(gdr._ffi.ptr(a), k, gdr._ffi.ptr(b), k, m, n, k, gdr._ffi.ptr(c), n,
                                        gdr._ffi.ptr(ws), ws.numel(), gdr._ffi.stream()), "gemm_nt_tc")
    torch.cuda.synchronize()
    return c


@pytest.mark.parametrize("m,n,k", [(128, 128, 32), (128, 128, 128), (256, 128, 512), (300, 260, 100),
                                   (1000, 256, 512), (5000, 256, 256), (77, 4, 8), (2048, 2048, 128)])
def test_gemm_nt_tc_is_fp32_grade(gdr, m, n, k):
    rng = np.random.default_rng(m * 7 + n * 3 + k)
    a = rng.standard_normal((m, k)).astype(np.float32)
    b = rng.standard_normal((n, k)).astype(np.float32)
    a[:, ::3] *= 37.0                                   # mixed magnitudes inside one dot product
    ref = a.astype(np.float64) @ b.astype(np.float64).T
    c = _gemm_nt_tc(gdr, torch.from_numpy(a).to(DEV), torch.from_numpy(b).to(DEV)).cpu().numpy()
    assert np.isfinite(c).all(), "unwritten or non-finite outputs"
    f32 = (torch.from_numpy(a) @ torch.from_numpy(b).T).numpy()     # the reference's arithmetic (torch CPU fp32)
    scale = np.abs(a.astype(np.float64)) @ np.abs(b.astype(np.float64)).T
    err_tc = np.abs(c - ref) / scale
    err_32 = np.abs(f32 - ref) / scale
    print(f"\n[tc gemm {m}x{n}x{k}] max err/scale: 3xTF32 {err_tc.max():.3e}  torch-cpu-fp32 {err_32.max():.3e}  "
          f"mean {err_tc.mean():.3e} vs {err_32.mean():.3e}")
    # fp32-grade: a few ulp of fp32 (2^-24 = 6e-8) relative to sum|a||b|; plain TF32 would sit at ~5e-4 here
    assert err_tc.max() <= max(4.0 * err_32.max(), 2e-6), "3xTF32 product is not fp32-grade"


def test_layer_forward_tc_matches_simt(gdr):
    L = gdr._ffi.lib()
    n, d = 3001, 256
    g = torch.Generator(device=DEV).manual_seed(0)
    p, t, res = (torch.randn(n, d, device=DEV, generator=g) for _ in range(3))
    w1, w2 = (torch.eye(d, device=DEV) + 0.05 * torch.randn(d, d, device=DEV, generator=g) for _ in range(2))
    b1, b2 = (0.1 * torch.randn(d, device=DEV, generator=g) for _ in range(2))
    out = {}
    for name in ("simt", "tc"):
        z, y = torch.empty(n, d, device=DEV), torch.empty(n, d, device=DEV)
        if name == "simt":
            gdr._ffi.check(L.gdr_gcn_layer_fwd_f32(*(gdr._ffi.ptr(v) for v in (p, t, w1, b1, w2, b2, res)), 0.3, n, d,
                                                   gdr._ffi.ptr(z), gdr._ffi.ptr(y), gdr._ffi.stream()))
        else:
            ws = gdr._ffi.workspace(L.gdr_gcn_layer_fwd_tc_workspace_bytes(n, d), p.device)
            gdr._ffi.check(L.gdr_gcn_layer_fwd_tc_f32(*(gdr._ffi.ptr(v) for v in (p, t, w1, b1, w2, b2, res)), 0.3, n,
                                                      d, gdr._ffi.ptr(z), gdr._ffi.ptr(y), gdr._ffi.ptr(ws),
                                                      ws.numel(), gdr._ffi.stream()))
        torch.cuda.synchronize()
        out[name] = (z.cpu().numpy(), y.cpu().numpy())
    # generic operands (|P·W| sums ~500 terms of mixed sign): two fp32 implementations differ by a few 1e-6 here, so
    # the yardstick is fp64 and the budget is the exact-fp32 kernel's own error
    z64 = (p.double() @ w1.double().T + b1.double() + t.double() @ w2.double().T + b2.double()).cpu().numpy()
    y64 = res.double().cpu().numpy() + 0.3 * np.where(z64 > 0, z64, np.expm1(z64))
    for i, (nm, ref) in enumerate((("Z", z64), ("Xnext", y64))):
        e_tc, e_simt = np.abs(out["tc"][i] - ref).max(), np.abs(out["simt"][i] - ref).max()
        print(f"\n[layer fwd {nm}] max abs err vs fp64: tc {e_tc:.3e} simt {e_simt:.3e}")
        assert e_tc <= max(2.0 * e_simt, 2e-6), f"{nm}: tc error {e_tc:.3e} vs exact-fp32 kernel {e_simt:.3e}"


def test_linear_backward_tc_matches_simt(gdr):
    L = gdr._ffi.lib()
    n, d = 2999, 128                                     # n not a multiple of 4: padded transposed operands
    g = torch.Generator(device=DEV).manual_seed(1)
    gz, a, e1, e2 = (torch.randn(n, d, device=DEV, generator=g) for _ in range(4))
    w = torch.eye(d, device=DEV) + 0.05 * torch.randn(d, d, device=DEV, generator=g)
    res = {}
    for name in ("simt", "tc"):
        c = torch.empty(n, d, device=DEV)
        dw, db = torch.ones(d, d, device=DEV), torch.ones(d, device=DEV)
        if name == "simt":
            gdr._ffi.check(L.gdr_linear_bwd_input_f32(gdr._ffi.ptr(gz), gdr._ffi.ptr(w), gdr._ffi.ptr(e1),
                                                      gdr._ffi.ptr(e2), n, d, gdr._ffi.ptr(c), gdr._ffi.stream()))
            ws = gdr._ffi.workspace(L.gdr_linear_bwd_weight_workspace_bytes(n, d), gz.device)
            gdr._ffi.check(L.gdr_linear_bwd_weight_f32(gdr._ffi.ptr(gz), gdr._ffi.ptr(a), n, d, 1, gdr._ffi.ptr(dw),
                                                       gdr._ffi.ptr(db), gdr._ffi.ptr(ws), ws.numel(),
                                                       gdr._ffi.stream()))
        else:
            ws = gdr._ffi.workspace(L.gdr_linear_bwd_input_tc_workspace_bytes(n, d), gz.device)
            gdr._ffi.check(L.gdr_linear_bwd_input_tc_f32(gdr._ffi.ptr(gz), gdr._ffi.ptr(w), gdr._ffi.ptr(e1),
                                                         gdr._ffi.ptr(e2), n, d, gdr._ffi.ptr(c), gdr._ffi.ptr(ws),
                                                         ws.numel(), gdr._ffi.stream()))
            ws = gdr._ffi.workspace(L.gdr_linear_bwd_weight_tc_workspace_bytes(n, d), gz.device)
            gdr._ffi.check(L.gdr_linear_bwd_weight_tc_f32(gdr._ffi.ptr(gz), gdr._ffi.ptr(a), n, d, 1, gdr._ffi.ptr(dw),
                                                          gdr._ffi.ptr(db), gdr._ffi.ptr(ws), ws.numel(),
                                                          gdr._ffi.stream()))
        torch.cuda.synchronize()
        res[name] = (c.cpu().numpy(), dw.cpu().numpy(), db.cpu().numpy())
    ref_dw = 1.0 + (gz.double().T @ a.double()).cpu().numpy()
    ref_dx = (gz.double() @ w.double() + e1.double() * e2.double()).cpu().numpy()
    e_tc, e_simt = np.abs(res["tc"][0] - ref_dx).max(), np.abs(res["simt"][0] - ref_dx).max()
    print(f"\n[dX] max abs err vs fp64: tc {e_tc:.3e} simt {e_simt:.3e}")
    assert e_tc <= max(2.0 * e_simt, 2e-6)
    # db: the tc path sums each column in fp64 inside the operand converter, the exact-fp32 kernel in fp32 slabs
    ref_db = 1.0 + gz.double().sum(0).cpu().numpy()
    e_tc, e_simt = np.abs(res["tc"][2] - ref_db).max(), np.abs(res["simt"][2] - ref_db).max()
    print(f"\n[db] max abs err vs fp64: tc {e_tc:.3e} simt {e_simt:.3e}")
    assert e_tc <= max(e_simt, 1e-6 * np.abs(ref_db).max())
    # dW sums 3k products per entry: compare both against fp64, the tc error must be of the same order
    e_tc, e_simt = np.abs(res["tc"][1] - ref_dw).max(), np.abs(res["simt"][1] - ref_dw).max()
    print(f"\n[dW] max abs err vs fp64: tc {e_tc:.3e} simt {e_simt:.3e}")
    assert e_tc <= max(4 * e_simt, 2e-5)


@pytest.mark.parametrize("n,d,acc", [(2999, 128, 1), (4097, 256, 0), (29959, 256, 0), (31, 128, 0), (1500, 64, 1),
                                     (3001, 384, 0)])
def test_weight_gradients_of_a_layer_tn_kernel(gdr, n, d, acc):
    """[dW1 | dW2] = Gᵀ·[A1 | A2], db1 = db2 = colsum(G) (modules/model.py:165,170 under autograd): the TN tcgen05
    kernel on raw operands (d % 128 == 0) and the transposing fallback (d = 64) against fp64."""
    L = gdr._ffi.lib()
    g = torch.Generator(device=DEV).manual_seed(n + d)
    gz, a1, a2 = (torch.randn(n, d, device=DEV, generator=g) for _ in range(3))
    gz[:, ::5] *= 19.0
    dw1, dw2 = (torch.full((d, d), 0.5, device=DEV) for _ in range(2))
    db1, db2 = (torch.full((d,), 0.5, device=DEV) for _ in range(2))
    ws = gdr._ffi.workspace(L.gdr_linear_bwd_weight2_tc_workspace_bytes(n, d), gz.device)
    gdr._ffi.check(L.gdr_linear_bwd_weight2_tc_f32(*(gdr._ffi.ptr(v) for v in (gz, a1, a2)), n, d, acc,
                                                   *(gdr._ffi.ptr(v) for v in (dw1, db1, dw2, db2)), gdr._ffi.ptr(ws),
                                                   ws.numel(), gdr._ffi.stream()))
    torch.cuda.synchronize()
    base = 0.5 if acc else 0.0
    gd = gz.double()
    for nm, got, ref, scale in (("dW1", dw1, gd.T @ a1.double(), gd.abs().T @ a1.double().abs()),
                                ("dW2", dw2, gd.T @ a2.double(), gd.abs().T @ a2.double().abs()),
                                ("db1", db1, gd.sum(0), gd.abs().sum(0)), ("db2", db2, gd.sum(0), gd.abs().sum(0))):
        err = ((got.double() - base - ref).abs() / scale).max().item()
        print(f"\n[{nm} n={n} d={d}] max err / sum|g||a|: {err:.3e}")
        assert err <= 2e-6, f"{nm}: {err:.3e}"
    assert torch.equal(db1, db2)


def test_training_steps_match_golden_with_tc_backend(gdr, monkeypatch):
    """The golden 3-Adam-step case of the reference (tests/golden/gcn_knn257.npz) with every dense contraction
    on the tensor cores."""
    from . import test_gpu_gcn
    monkeypatch.setattr(gdr.model, "GEMM_BACKEND", "tc")
    for case in ("gcn_knn257.npz", "gcn_knn600_l3.npz", "gcn_edge300_d256.npz"):
        test_gpu_gcn.test_training_steps_match_reference(gdr, case)
    test_gpu_gcn.test_layer_forward_matches_reference(gdr)
    test_gpu_gcn.test_input_gradient_and_oracle_autograd(gdr)


def test_gss_loss_tc_matches_simt(gdr, monkeypatch):
    n, d, b = 1500, 128, 777                               # b not a multiple of 4: padded B x B buffers
    g = torch.Generator(device=DEV).manual_seed(3)
    h = torch.nn.functional.normalize(torch.randn(n, d, device=DEV, generator=g), dim=1)
    idx = torch.randperm(n, device=DEV, generator=g)[:b]
    out = {}
    for name in ("simt", "tc"):
        monkeypatch.setattr(gdr.model, "GEMM_BACKEND", name)
        e = h.clone().requires_grad_(True)
        loss = gdr.GSS_loss(1.0).gss_loss(embs=e, beta=0.2, index=idx)
        loss.backward()
        torch.cuda.synchronize()
        out[name] = (loss.item(), e.grad.cpu().numpy())
    assert_close(out["tc"][0], out["simt"][0], what="loss tc vs simt")
    assert_close(out["tc"][1], out["simt"][1], what="dLoss/dE tc vs simt")
    # and against torch autograd on the CPU (the reference's arithmetic, modules/model.py:214-221)
    e = h.cpu().clone().requires_grad_(True)
    ee = e[idx.cpu()]
    gm = torch.mm(ee, ee.t())
    ref = torch.mean(-0.5 * (torch.relu(gm) - 0.2) ** 2)
    ref.backward()
    assert_close(out["tc"][0], ref.item(), what="loss tc vs torch cpu")
    assert_close(out["tc"][1], e.grad.numpy(), what="dLoss/dE tc vs torch cpu")


@pytest.mark.parametrize("n,d,rows", [(1000, 128, 256), (1333, 64, 128), (600, 256, 384)])
def test_gss_loss_full_batch_in_row_panels(gdr, monkeypatch, n, d, rows):
    """index=None (train.py --batch-size 0): the B x B Gram is processed in row panels and never stored
    (csrc/loss.cu).  Several panels (forced small) must give the loss and gradient of modules/model.py:214-221."""
    g = torch.Generator(device=DEV).manual_seed(n)
    h = torch.nn.functional.normalize(torch.randn(n, d, device=DEV, generator=g), dim=1)
    res = {}
    for name, env in (("one_panel", None), ("panels", str(rows))):
        if env is None:
            monkeypatch.delenv("GDR_GSS_PANEL_ROWS", raising=False)
        else:
            monkeypatch.setenv("GDR_GSS_PANEL_ROWS", env)
        e = h.clone().requires_grad_(True)
        loss = gdr.GSS_loss(1.0).gss_loss(embs=e, beta=0.15)
        loss.backward()
        torch.cuda.synchronize()
        res[name] = (loss.item(), e.grad.cpu().numpy())
    e = h.cpu().double().clone().requires_grad_(True)
    ref = torch.mean(-0.5 * (torch.relu(torch.mm(e, e.t())) - 0.15) ** 2)
    ref.backward()
    for name in res:
        assert_close(res[name][0], ref.item(), what=f"loss ({name}) vs torch fp64")
        assert_close(res[name][1], e.grad.numpy(), what=f"dLoss/dE ({name}) vs torch fp64")
    assert_close(res["panels"][1], res["one_panel"][1], what="panels vs one panel")


def test_gss_loss_full_batch_workspace_is_bounded(gdr):
    """B = 2^20, d = 256 (RMAT-20): the workspace is a few GB of panels, not the 4 TB of B x B."""
    L = gdr._ffi.lib()
    b, d = 1 << 20, 256
    ws = L.gdr_gss_loss_tc_workspace_bytes(b, d)
    assert ws < 12 * (1 << 30), ws
    assert ws < b * b * 4 // 100


@pytest.mark.parametrize("n,d,q", [(257, 128, 98.0), (600, 128, 98.0), (300, 256, 50.0), (130, 12, 0.0), (130, 12, 100.0)])
def test_gram_percentile_tc(gdr, monkeypatch, n, d, q):
    from . import test_gpu_gcn
    monkeypatch.setattr(gdr.model, "GEMM_BACKEND", "tc")
    test_gpu_gcn.test_gram_percentile_vs_numpy(gdr, n, d, q)


def test_beta_selection_tc(gdr, monkeypatch):
    from . import test_gpu_gcn
    monkeypatch.setattr(gdr.model, "GEMM_BACKEND", "tc")
    for case in ("gcn_knn257.npz", "gcn_knn600_l3.npz", "gcn_edge300_d256.npz"):
        test_gpu_gcn.test_beta_selection_matches_reference(gdr, case)


def test_fused_adam_matches_torch_adam(gdr):
    """gdr.optim.Adam (one launch for all tensors) against torch.optim.Adam on the same gradients, 6 steps;
    state_dict layouts are interchangeable."""
    g = torch.Generator(device=DEV).manual_seed(5)
    shapes = [(256, 256), (256,), (256, 256), (256,)]
    p_ref = [torch.randn(*s, device=DEV, generator=g).requires_grad_(True) for s in shapes]
    p_our = [p.detach().clone().requires_grad_(True) for p in p_ref]
    o_ref = torch.optim.Adam(p_ref, lr=3e-4, weight_decay=0)
    o_our = gdr.optim.Adam(p_our, lr=3e-4, weight_decay=0)
    for step in range(6):
        for a, b in zip(p_ref, p_our):
            gr = torch.randn(a.shape, device=DEV, generator=g) * (10.0 ** (step - 3))
            a.grad, b.grad = gr.clone(), gr.clone()
        o_ref.step()
        o_our.step()
    torch.cuda.synchronize()
    for a, b in zip(p_ref, p_our):
        assert_close(b.detach().cpu().numpy(), a.detach().cpu().numpy(), rtol=1e-6, atol=1e-7, what="param")
    sd_ref, sd_our = o_ref.state_dict(), o_our.state_dict()
    assert sd_ref["state"].keys() == sd_our["state"].keys()
    for k in sd_ref["state"]:
        assert set(sd_ref["state"][k]) == set(sd_our["state"][k]) == {"step", "exp_avg", "exp_avg_sq"}
        assert float(sd_ref["state"][k]["step"]) == float(sd_our["state"][k]["step"]) == 6.0
        assert_close(sd_our["state"][k]["exp_avg_sq"].cpu().numpy(), sd_ref["state"][k]["exp_avg_sq"].cpu().numpy(),
                     rtol=1e-6, atol=0, what="exp_avg_sq")
    o_ref.load_state_dict(sd_our)                          # the reference optimizer accepts our state
