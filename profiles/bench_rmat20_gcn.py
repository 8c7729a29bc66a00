"""BASELINE configs[3]: GCN training iteration (2-layer GSS GCN, d 256, batch 2048 GSS loss, backward, Adam) on a
synthetic R-MAT graph with 1M nodes / ~32M edges on one B200.
    python profiles/bench_rmat20_gcn.py [scale] [iters]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import gdr_b200 as gdr  # noqa: E402

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 20
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda", 0)
d, b = 256, 2048
csr = bench.build_rmat_csr(scale, 32, dev)
n = csr.n_rows
g = torch.Generator(device=dev).manual_seed(2)
x = torch.randn(n, d, device=dev, generator=g)
np.random.seed(1)
torch.manual_seed(1)
model = gdr.ResidualGraphConvolutionalNetwork(b, n, num_layers=2, hidden_units=d, layer_decay=0.3).to(dev)
opt = gdr.optim.Adam(model.parameters(), lr=3e-4, weight_decay=0)
loss_fcn = gdr.GSS_loss(1.0).gss_loss


def iteration():
    idx = torch.randint(0, n, (b,), device=dev, generator=g)
    h = model(x=x, adj=csr)
    loss = loss_fcn(embs=h, beta=0.1, index=idx)
    opt.zero_grad()
    loss.backward()
    opt.step()
    return loss


for _ in range(2):
    iteration()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    loss = iteration()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(json.dumps({"workload": "rmat_gcn: 2-layer GSS GCN training iteration, R-MAT scale %d edge-factor 32, d %d, batch %d "
                              "(BASELINE configs[3])" % (scale, d, b), "nodes": n, "nnz": csr.nnz,
                  "ms_per_iteration": ms, "spmm_launches_per_iteration": 6,
                  "gedges_per_s_through_the_spmms": 6 * csr.nnz / (ms * 1e-3) / 1e9,
                  "loss": float(loss.item()), "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9}))
