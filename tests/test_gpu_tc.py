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
    assert err_tc.max() <= max(4.0 * err_32.max(), 3e-7), "3xTF32 product is not fp32-grade"


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
    assert_close(out["tc"][0], out["simt"][0], what="Z tc vs simt")
    assert_close(out["tc"][1], out["simt"][1], what="Xnext tc vs simt")


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
    assert_close(res["tc"][0], res["simt"][0], what="dX tc vs simt")
    assert_close(res["tc"][2], res["simt"][2], what="db tc vs simt")
    # dW sums 3k products per entry: compare both against fp64, the tc error must be of the same order
    e_tc, e_simt = np.abs(res["tc"][1] - ref_dw).max(), np.abs(res["simt"][1] - ref_dw).max()
    print(f"\n[dW] max abs err vs fp64: tc {e_tc:.3e} simt {e_simt:.3e}")
    assert e_tc <= max(4 * e_simt, 2e-5)


def test_training_steps_match_golden_with_tc_backend(gdr, monkeypatch):
    """The golden 3-Adam-step case of the reference (tests/golden/gcn_knn257.npz) with every dense contraction
    on the tensor cores."""
    from . import test_gpu_gcn
    monkeypatch.setattr(gdr.model, "GEMM_BACKEND", "tc")
    for case in ("gcn_knn257.npz", "gcn_knn600_l3.npz", "gcn_edge300_d256.npz"):
        test_gpu_gcn.test_training_steps_match_reference(gdr, case)
    test_gpu_gcn.test_layer_forward_matches_reference(gdr)
    test_gpu_gcn.test_input_gradient_and_oracle_autograd(gdr)
