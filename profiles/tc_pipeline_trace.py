"""Pipeline trace + timing of the tcgen05 GEMM kernels (diagnostic; `gdr_debug_tc_trace`, csrc/gemm_tc.cu).

    python profiles/tc_pipeline_trace.py [out.json]

Times the layer-forward GEMM (EPI_LAYER, A = [P | T] raw) and the plain NT GEMM (EPI_STORE) at the MSI shape with
CUDA events, then records one launch of each with CTA 0's clock64() stamps at every hand-off of the producer /
converter / MMA / epilogue roles and prints the per-k-block intervals.  Needs a B200; writes JSON next to the log.
"""
import ctypes
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import gdr_b200 as gdr  # noqa: E402

DEV = "cuda:0"
F = gdr._ffi
L = F.lib()


def time_it(fn, reps=20):
    for _ in range(3):
        fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return float(np.median(ts)), float(np.min(ts))


def trace(fn):
    buf = torch.zeros(4 * 128 * 4, dtype=torch.int64, device=DEV)
    L.gdr_debug_tc_trace.argtypes, L.gdr_debug_tc_trace.restype = [ctypes.c_void_p], ctypes.c_int   # not in gdr.h
    F.check(L.gdr_debug_tc_trace(F.ptr(buf)))
    fn()
    torch.cuda.synchronize()
    F.check(L.gdr_debug_tc_trace(None))
    return buf.cpu().numpy().reshape(4, 128, 4)


def summarize(tr, name):
    t0 = tr[tr > 0].min()
    r = np.where(tr > 0, tr - t0, -1)
    nk = int((r[1, :, 2] >= 0).sum())
    print(f"--- {name}: {nk} k-blocks traced on CTA 0, span {r.max()} clk")
    print("it | prod: aempty  bempty  issued | conv: afull  cvt-done  tempty  stored | mma: bfull  tfull  committed")
    for it in list(range(min(nk, 24))) + list(range(max(24, nk - 6), nk)):
        print(f"{it:3d} | {r[0, it, 0]:7d} {r[0, it, 1]:7d} {r[0, it, 2]:7d} | {r[2, it, 0]:7d} {r[2, it, 1]:7d} "
              f"{r[2, it, 2]:7d} {r[2, it, 3]:7d} | {r[1, it, 0]:7d} {r[1, it, 1]:7d} {r[1, it, 2]:7d}")
    nt = int((r[3, :, 2] >= 0).sum())
    for tl in range(nt):
        print(f"epilogue tile {tl}: acc_full {r[3, tl, 0]}  acc released {r[3, tl, 1]}  done {r[3, tl, 2]}")
    mma = r[1, :nk, 2]
    if nk > 8:
        print(f"MMA commit interval: median {np.median(np.diff(mma)):.0f} clk, mean {np.diff(mma).mean():.0f} clk")
    return {"k_blocks": nk, "span_clk": int(r.max()), "mma_commit_interval_clk_median": float(np.median(np.diff(mma))) if nk > 8 else None,
            "epilogue": [[int(x) for x in r[3, tl, :3]] for tl in range(nt)]}


def main():
    out = {}
    n, d = 29959, 256
    g = torch.Generator(device=DEV).manual_seed(0)
    p, t, res = (torch.randn(n, d, device=DEV, generator=g) for _ in range(3))
    w1, w2 = (torch.eye(d, device=DEV) + 0.05 * torch.randn(d, d, device=DEV, generator=g) for _ in range(2))
    b1, b2 = (0.1 * torch.randn(d, device=DEV, generator=g) for _ in range(2))
    z, y = torch.empty(n, d, device=DEV), torch.empty(n, d, device=DEV)
    ws = F.workspace(L.gdr_gcn_layer_fwd_tc_workspace_bytes(n, d), p.device)

    def layer(resid=res):
        F.check(L.gdr_gcn_layer_fwd_tc_f32(*(F.ptr(v) for v in (p, t, w1, b1, w2, b2)), F.ptr(resid) if resid is not None else None,
                                           0.3, n, d, F.ptr(z), F.ptr(y), F.ptr(ws), ws.numel(), F.stream()))

    for nm, fn in (("layer_fwd_29959x256x512", layer), ("layer_fwd_no_residual", lambda: layer(None))):
        med, mn = time_it(fn)
        print(f"{nm}: median {med:.1f} us  min {mn:.1f} us (includes the 2-launch weight split)")
        out[nm] = {"median_us": med, "min_us": mn}
    out["layer_trace"] = summarize(trace(layer), "layer forward")

    for (m, nn, k) in ((29959, 256, 512), (29959, 256, 256), (29959, 128, 512), (4096, 4096, 256), (18944, 256, 512)):
        a = torch.randn(m, k, device=DEV, generator=g)
        b = torch.randn(nn, k, device=DEV, generator=g)
        c = torch.empty(m, nn, device=DEV)
        ws2 = F.workspace(L.gdr_gemm_nt_tc_workspace_bytes(m, nn, k), a.device)

        def gemm():
            F.check(L.gdr_gemm_nt_tc_f32(F.ptr(a), k, F.ptr(b), k, m, nn, k, F.ptr(c), nn, F.ptr(ws2), ws2.numel(),
                                         F.stream()))
        med, mn = time_it(gemm)
        tf = 2.0 * m * nn * k / (med * 1e-6) / 1e12
        print(f"gemm_nt {m}x{nn}x{k}: median {med:.1f} us  min {mn:.1f} us  ({tf:.1f} fp32-equivalent TFLOP/s, "
              f"{3 * tf:.1f} TF32 TFLOP/s issued)")
        out[f"gemm_nt_{m}x{nn}x{k}"] = {"median_us": med, "min_us": mn, "fp32_equiv_tflops": tf}
        if (m, nn, k) == (29959, 256, 512):
            out["gemm_trace"] = summarize(trace(gemm), "plain NT GEMM 29959x256x512")
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
