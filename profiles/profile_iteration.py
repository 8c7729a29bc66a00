"""ncu target: ONE training iteration of the bench workload (msi_gcn) or a few SpMM launches (rmat),
bracketed by cudaProfilerStart/Stop so that `ncu --profile-from-start off` sees only that region.

    python profiles/profile_iteration.py msi            # one fwd+loss+bwd+Adam iteration at MSI scale
    python profiles/profile_iteration.py rmat 22 128    # 3 SpMM launches on R-MAT scale 22, d 128
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import gdr_b200 as gdr  # noqa: E402
from gdr_b200 import trainer as trainer_mod  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "msi"
dev = torch.device("cuda", 0)
if mode == "msi":
    a, feats = bench.msi_inputs()
    adj = gdr.convert_sparse_matrix_to_sparse_tensor(gdr.preprocess_graph(a))
    np.random.seed(1)
    torch.manual_seed(1)
    tr = trainer_mod.GCNTrainer(feats, adj, hidden_units=feats.shape[1], num_layers=2, layer_decay=0.3, lr=3e-4,
                                batch_size=2048, beta_percentile=98.0, hoist_layer1=False)   # the bench's `value` path
    batches = [b.to(dev) for b in bench.epoch_batches(feats.shape[0], 2048, 1, 1)[0]]
    for b in batches[:3]:
        tr.iteration(b)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    tr.iteration(batches[3])
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
else:
    scale, d = int(sys.argv[2]), int(sys.argv[3])
    csr = bench.build_rmat_csr(scale, 32, dev)
    x = torch.randn(csr.n_rows, d, device=dev)
    y = torch.empty_like(x)
    for _ in range(2):
        csr.spmm(x, out=y)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    for _ in range(3):
        csr.spmm(x, out=y)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
print("profile target done")
