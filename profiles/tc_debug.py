"""First-light check of the tcgen05 GEMM: tiny shapes, prints what kind of wrong (if any) the output is.
    timeout 120 python profiles/tc_debug.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gdr_b200 as gdr  # noqa: E402

L = gdr._ffi.lib()
dev = torch.device("cuda", 0)


def run(m, n, k, seed=0, reps=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    a = torch.randn(m, k, device=dev, generator=g)
    b = torch.randn(n, k, device=dev, generator=g)
    c = torch.full((m, n), float("nan"), device=dev)
    ws = gdr._ffi.workspace(L.gdr_gemm_nt_tc_workspace_bytes(m, n, k), dev)
    rc = L.gdr_gemm_nt_tc_f32(a.data_ptr(), k, b.data_ptr(), k, m, n, k, c.data_ptr(), n, ws.data_ptr(), ws.numel(),
                              torch.cuda.current_stream().cuda_stream)
    if rc != 0:
        print("rc", rc, L.gdr_last_error_string().decode())
        return
    torch.cuda.synchronize()
    ref = (a.double() @ b.double().T)
    err = (c.double() - ref).abs()
    scale = a.double().abs() @ b.double().abs().T
    f32 = (a @ b.T).double()
    print(f"{m}x{n}x{k}: nan {int(torch.isnan(c).sum())}  max|err| {float(err.nan_to_num(1e9).max()):.3e}  "
          f"max err/scale {float((err / scale).nan_to_num(1e9).max()):.3e}  (cuBLAS fp32: "
          f"{float(((f32 - ref).abs() / scale).max()):.3e})")
    if float(err.nan_to_num(1e9).max()) > 1e-2:
        # classify: tf32-grade? transposed? row/col permuted?
        tf = (a.double() @ b.double().T)
        print("  c[0,:6]  ", c[0, :6].tolist())
        print("  ref[0,:6]", ref[0, :6].tolist())
        print("  c[:6,0]  ", c[:6, 0].tolist())
        print("  ref[:6,0]", ref[:6, 0].tolist())
        ratio = (c.double() / ref).flatten()
        print("  median c/ref", float(ratio.nan_to_num(0).median()))
    if reps:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            L.gdr_gemm_nt_tc_f32(a.data_ptr(), k, b.data_ptr(), k, m, n, k, c.data_ptr(), n, ws.data_ptr(), ws.numel(),
                                 torch.cuda.current_stream().cuda_stream)
        e0.record()
        for _ in range(reps):
            L.gdr_gemm_nt_tc_f32(a.data_ptr(), k, b.data_ptr(), k, m, n, k, c.data_ptr(), n, ws.data_ptr(), ws.numel(),
                                 torch.cuda.current_stream().cuda_stream)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print(f"  {ms * 1e3:.1f} us per call (split + gemm), {2.0 * m * n * k / ms / 1e9:.1f} TFLOP/s fp32-equivalent")


for shape in [(128, 128, 32), (128, 128, 64), (128, 128, 512), (256, 256, 128), (300, 260, 100)]:
    run(*shape)
run(29959, 256, 512, reps=20)
run(29959, 256, 256, reps=20)
run(2048, 2048, 256, reps=20)
print("tc_debug done")
