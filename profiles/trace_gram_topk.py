"""Pipeline trace of gram_topk_tc_kernel (diagnostic; gdr_debug_tc_trace): CTA 0's clock64() stamps in steady state
(k-blocks 512.., tiles 64..) at N x d from the command line."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gdr_b200 as gdr
F = gdr._ffi; L = F.lib()
n, d = int(sys.argv[1]), int(sys.argv[2])
g = torch.Generator(device="cuda").manual_seed(2)
x = torch.randn(n, d, device="cuda", dtype=torch.float64, generator=g)
gdr.helper.gram_topk(x, 5, backend="tc")
buf = torch.zeros(4 * 128 * 4, dtype=torch.int64, device="cuda")
F.check(L.gdr_debug_tc_trace(F.ptr(buf)))
gdr.helper.gram_topk(x, 5, backend="tc")
torch.cuda.synchronize()
F.check(L.gdr_debug_tc_trace(None))
tr = buf.cpu().numpy().reshape(4, 128, 4)
t0 = tr[tr > 0].min()
r = np.where(tr > 0, tr - t0, -1)
print("it | prod: wait-empty got-empty | mma: wait-full got-full committed (acc_empty at tile start)")
for it in range(0, 64):
    print(f"{it:3d} | {r[0,it,0]:7d} {r[0,it,1]:7d} | {r[1,it,0]:7d} {r[1,it,1]:7d} {r[1,it,2]:7d} {r[1,it,3]:7d}")
print("tile | epi: wait-acc got-acc released done")
for tl in range(0, 16):
    print(f"{tl:3d} | {r[3,tl,0]:7d} {r[3,tl,1]:7d} {r[3,tl,2]:7d} {r[3,tl,3]:7d}")
c = r[1, :, 2]; c = c[c >= 0]
print("MMA commit interval per k-block: median", np.median(np.diff(c)), "mean", np.diff(c).mean())
e = r[3, :, 3] - r[3, :, 1]; print("epilogue busy per tile (got-acc -> done): median", np.median(e[r[3,:,3] >= 0]))
e = r[3, :, 1] - r[3, :, 0]; print("epilogue wait for acc_full per tile: median", np.median(e[r[3,:,1] >= 0]))
