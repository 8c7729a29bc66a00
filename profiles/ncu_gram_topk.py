"""ncu target: one tensor-core kNN selection (gdr_gram_topk_tc_f64) at N x d given on the command line."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gdr_b200 as gdr
n, d = int(sys.argv[1]), int(sys.argv[2])
g = torch.Generator(device="cuda").manual_seed(2)
x = torch.randn(n, d, device="cuda", dtype=torch.float64, generator=g)
gdr.helper.gram_topk(x, 5, backend="tc")
torch.cuda.synchronize()
torch.cuda.profiler.start()
gdr.helper.gram_topk(x, 5, backend="tc")
torch.cuda.synchronize()
torch.cuda.profiler.stop()
