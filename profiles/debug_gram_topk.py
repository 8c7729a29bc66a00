"""Diagnostic: compare the tensor-core kNN selection with a direct fp64 product on ALL rows (in row chunks) and
histogram the mismatching rows by position in the 128-row tile and by work-item ordinal."""
import sys, os, collections
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gdr_b200 as gdr
n = int(sys.argv[1]); d = int(sys.argv[2]); k = int(sys.argv[3]) if len(sys.argv) > 3 else 5
g = torch.Generator(device="cuda").manual_seed(2)
x = torch.randn(n, d, device="cuda", dtype=torch.float64, generator=g)
x *= torch.empty(n, 1, device="cuda", dtype=torch.float64).uniform_(0.5, 3.0, generator=g)
idx, val = gdr.helper.gram_topk(x, k, backend="tc")
nfb = int(gdr.helper.last_gram_topk_info["fallback_rows"].item())
bad_rows = []
for r0 in range(0, n, 8192):
    tv, ti = torch.topk(x[r0:r0 + 8192] @ x.T, k, dim=1)
    bad = ~(torch.sort(ti, 1).values == torch.sort(idx[r0:r0 + 8192].long(), 1).values).all(1)
    bad_rows += (bad.nonzero().flatten() + r0).tolist()
print("n", n, "d", d, "fallback", nfb, "bad", len(bad_rows))
print("by r%128:", sorted(collections.Counter(r % 128 for r in bad_rows).items()))
print("by ordinal:", sorted(collections.Counter((r // 128) // 148 for r in bad_rows).items()))
print("by cta:", sorted(collections.Counter((r // 128) % 148 for r in bad_rows).items())[:40])
