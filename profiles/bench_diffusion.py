"""BASELINE configs[2]: MSI diffusion profiles (random walk with restart) for ALL drugs and indications in one
batched device solve, on the MSI-shaped synthetic interactome (29,959 nodes / 957,456 directed edges / 2,501
seeds), next to the reference algorithm (oracle/rwr_oracle.py = multiscale/diff_prof/diffusion_profiles.py:29-90
restated) timed on a few seeds on one host core.
    python profiles/bench_diffusion.py [n_cpu_seeds]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gdr_b200 as gdr  # noqa: E402
from oracle import rwr_oracle  # noqa: E402
from tests.test_gpu_rwr import SynthMSI  # noqa: E402

n_cpu = int(sys.argv[1]) if len(sys.argv) > 1 else 8
g = gdr.synth.msi_synth(seed=20200430, scale=1.0)
msi = SynthMSI(g, gdr.synth)
dp = gdr.DiffusionProfiles(alpha=gdr.synth.RWR_ALPHA, max_iter=gdr.synth.RWR_MAX_ITER, tol=gdr.synth.RWR_TOL,
                           weights=None, num_cores=4, save_load_file_path=None)
t0 = time.perf_counter()
dp.get_initial_M(msi)
t_m = time.perf_counter() - t0
seeds = msi.drugs_in_graph + msi.indications_in_graph
dp.solve(msi, seeds[:64])                                   # builds the device problem, warms the kernels up
torch.cuda.synchronize()
t0 = time.perf_counter()
x, iters = dp.solve(msi, seeds)                             # includes the D2H copy of all profiles (600 MB)
torch.cuda.synchronize()
t_gpu = time.perf_counter() - t0
nnz = int(dp.initial_M.nnz)
m0 = rwr_oracle.initial_matrix(g["n"], g["src"], g["dst"], g["weight"])
sinks = [int(s) for s in g["seeds"]]
t0 = time.perf_counter()
worst = 0.0
for nm in seeds[:n_cpu]:
    ox, _ = rwr_oracle.diffusion_profile(m0, msi.node2idx[nm], g["seed_proteins"], sinks, gdr.synth.RWR_ALPHA,
                                         gdr.synth.RWR_TOL, gdr.synth.RWR_MAX_ITER)
    worst = max(worst, float(np.abs(x[seeds.index(nm)] - ox).max() / np.abs(ox).max()))
t_cpu = (time.perf_counter() - t0) / max(n_cpu, 1)
print(json.dumps({
    "workload": "msi_diffusion_profiles: all drugs + indications, batched fp64 power iteration (BASELINE configs[2])",
    "nodes": g["n"], "nnz": nnz, "seeds": len(seeds), "seconds_all_seeds_incl_d2h": t_gpu,
    "seeds_per_s": len(seeds) / t_gpu, "iterations_min_max": [int(iters.min()), int(iters.max())],
    "effective_gedges_per_s": float(iters.sum()) * nnz / t_gpu / 1e9, "host_matrix_build_s": t_m,
    "cpu_reference_algorithm_s_per_seed_1_core": t_cpu, "cpu_seeds_timed": n_cpu,
    "speedup_vs_1_core": t_cpu * len(seeds) / t_gpu, "max_rel_diff_vs_oracle_on_timed_seeds": worst}))
