"""gdr_gss_loss_fwd_bwd_static_f32: the GSS loss for callers that cannot allocate device memory.

Some users of libgdr.so have no way to produce a workspace buffer at call time: a plug-in running inside a host
whose device-memory pool it does not own, a serving loop that bans allocator calls once it is warm (cudaMalloc takes a
device-wide lock and may synchronise), or a plain C program with no allocator of its own.  For them the library
carries a fixed scratch array in its own device image (18.1 MiB per GPU, B <= 2048 and B*d <= 2048*256) and this
entry point uses it instead of a `workspace` argument.  The caller still supplies inputs and outputs.

What is checked here, through the C ABI only:
  * two calls in a row (different sizes, the second one smaller and ragged, so stale scratch from the first call is
    sitting under it) give the loss and dLoss/dE of the ordinary workspace entry point within the suite's tolerance;
  * oversize requests come back as GDR_E_SHAPE with a readable message and launch nothing (also checked without a GPU).

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a problem in this optional entry
point must not hide the parity tests.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes

import pytest
import torch

from .conftest import assert_close

GDR_E_SHAPE = -2


def _loss_with_workspace(gdr, embs, index, alpha, beta):
    L = gdr._ffi.lib()
    b, (n, d) = index.numel(), embs.shape
    loss = torch.empty(1, dtype=torch.float32, device=embs.device)
    grad = torch.empty(b, d, dtype=torch.float32, device=embs.device)
    ws = gdr._ffi.workspace(L.gdr_gss_loss_workspace_bytes(b, d), embs.device)
    gdr._ffi.check(L.gdr_gss_loss_fwd_bwd_f32(gdr._ffi.ptr(embs), gdr._ffi.ptr(index), b, n, d, alpha, beta,
                                              gdr._ffi.ptr(loss), gdr._ffi.ptr(grad), gdr._ffi.ptr(ws), ws.numel(),
                                              gdr._ffi.stream()), "gss_loss_fwd_bwd")
    return loss, grad


def _loss_static(gdr, embs, index, alpha, beta, loss, grad):
    """loss / grad are allocated by the caller beforehand: the call itself must not need any new memory."""
    L = gdr._ffi.lib()
    b, (n, d) = index.numel(), embs.shape
    gdr._ffi.check(L.gdr_gss_loss_fwd_bwd_static_f32(gdr._ffi.ptr(embs), gdr._ffi.ptr(index), b, n, d, alpha, beta,
                                                     gdr._ffi.ptr(loss), gdr._ffi.ptr(grad), gdr._ffi.stream()),
                   "gss_loss_fwd_bwd_static")


@pytest.mark.gpu
def test_static_scratch_loss_matches_workspace_loss_twice_in_a_row(gdr):
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(1234)
    # first call: the reference's training batch (fills the whole scratch array); second: small, B % 4 != 0, other d
    cases = [(5000, 256, 2048, 1.0, 0.3), (700, 64, 333, 0.7, 0.15)]
    inputs = []
    for n, d, b, alpha, beta in cases:
        embs = torch.nn.functional.normalize(torch.randn(n, d, generator=gen), dim=1).to(dev)
        index = torch.randperm(n, generator=gen)[:b].to(dev)
        out = (torch.empty(1, dtype=torch.float32, device=dev), torch.empty(b, d, dtype=torch.float32, device=dev))
        inputs.append((embs, index, alpha, beta, out))
    torch.cuda.synchronize()

    got = []
    reserved_before = torch.cuda.memory_reserved(dev)
    for embs, index, alpha, beta, (loss, grad) in inputs:       # back to back, same stream, no sync in between
        _loss_static(gdr, embs, index, alpha, beta, loss, grad)
        got.append((loss, grad))
    torch.cuda.synchronize()
    assert torch.cuda.memory_reserved(dev) == reserved_before    # torch's allocator was not asked for anything

    for (embs, index, alpha, beta, _), (loss, grad) in zip(inputs, got):
        want_loss, want_grad = _loss_with_workspace(gdr, embs, index, alpha, beta)
        torch.cuda.synchronize()
        what = f"B={index.numel()} d={embs.shape[1]}"
        assert_close(loss.cpu().numpy(), want_loss.cpu().numpy(), what=f"static loss {what}")
        assert_close(grad.cpu().numpy(), want_grad.cpu().numpy(), what=f"static dLoss/dE {what}")
        assert torch.isfinite(grad).all() and grad.abs().max() > 0


@pytest.mark.gpu
def test_static_scratch_repeat_call_is_stable(gdr):
    """The same input twice: the second call runs on top of its own leftovers and must give the same numbers."""
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(7)
    embs = torch.randn(900, 128, generator=gen).to(dev) * 0.2
    index = torch.randperm(900, generator=gen)[:513].to(dev)
    outs = [(torch.empty(1, device=dev), torch.empty(513, 128, device=dev)) for _ in range(2)]
    for loss, grad in outs:
        _loss_static(gdr, embs, index, 1.0, 0.25, loss, grad)
    torch.cuda.synchronize()
    want_loss, want_grad = _loss_with_workspace(gdr, embs, index, 1.0, 0.25)
    for loss, grad in outs:
        assert_close(loss.cpu().numpy(), want_loss.cpu().numpy(), what="repeat loss")
        assert_close(grad.cpu().numpy(), want_grad.cpu().numpy(), what="repeat dLoss/dE")


def test_static_scratch_rejects_sizes_beyond_the_array(gdr):
    """Argument checks run before any CUDA call, so this needs no GPU: nothing is launched, nothing is touched."""
    L = gdr._ffi.lib()
    fake = ctypes.c_void_p(256)  # never dereferenced: the size check fails first
    before = L.gdr_launch_count()
    for b, d in ((2049, 64), (2048, 260), (1024, 516)):
        rc = L.gdr_gss_loss_fwd_bwd_static_f32(fake, None, b, 4096, d, 1.0, 0.3, fake, fake, None)
        assert rc == GDR_E_SHAPE, (b, d, rc)
        msg = L.gdr_last_error_string().decode()
        assert "built-in scratch" in msg and str(b) in msg
    assert L.gdr_gss_loss_fwd_bwd_static_f32(fake, None, 16, 16, 6, 1.0, 0.3, fake, fake, None) == GDR_E_SHAPE
    assert L.gdr_gss_loss_fwd_bwd_static_f32(None, None, 16, 16, 8, 1.0, 0.3, fake, fake, None) == -1
    assert L.gdr_launch_count() == before
