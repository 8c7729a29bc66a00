#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 GCN-propagation path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload ...]

Workloads (BASELINE.json `metric`: "GCN epochs/s on MSI graph; SpMM GEdges/s + HBM GB/s at 1/2/4/8 B200"):

  msi_gcn    (default at N=1; BASELINE configs[1]) 2-layer GSS GCN, hidden 256, on the MSI-shaped weighted
             graph (29,959 nodes / 478,728 undirected edges -> nnz(Â) = 987,415), batch 2048, Adam lr 3e-4,
             train.sh hyper-parameters.  One step = one epoch = ceil(N/B) = 15 iterations of
             (full-graph forward + batch GSS loss + backward + Adam), train.py:155-184.  metric = epochs/s.
             The same line carries `spmm_rmat`: the HBM-bound SpMM number on an R-MAT graph at 1 GPU.
  rmat_spmm  (default at N>1; BASELINE configs[4]) SpMM Y = Â·X on an R-MAT graph (scale 24, edge factor 32,
             d=128), L-shaped partition over the ranks (gcn-drug-repurposing_b200/dist.py): only the referenced
             X rows / the partial Y rows travel, as direct stores into peer symmetric memory over NVLink
             (--transport nccl: all_to_all of packed rows; --hubs K: NVSwitch multicast for the hub ranks;
             --partition rows: plain row blocks).  One step = one SpMM over the whole graph, exchanges included.
             metric = GEdges/s (strong scaling); the line carries the same product's 1-GPU time.

One JSON line on stdout (rank 0).  See DESIGN.md §Measurement for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MSI_D = 256            # BASELINE configs[1]: "hidden 256"
MSI_BATCH = 2048       # train.sh
MSI_LAYERS = 2
MSI_DECAY = 0.3        # train.py:39 default (train.sh does not override)
MSI_LR = 3e-4
MSI_BETA_PCT = 98.0
L2_FLUSH_BYTES = 256 << 20


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", choices=["ours", "reference"], default="ours")
    p.add_argument("--workload", choices=["auto", "msi_gcn", "rmat_spmm"], default="auto")
    p.add_argument("--rmat-scale", type=int, default=24)
    p.add_argument("--rmat-edge-factor", type=int, default=32)
    p.add_argument("--rmat-d", type=int, default=128)
    p.add_argument("--no-rmat", action="store_true", help="skip the secondary R-MAT SpMM measurement at N=1")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-overlap", action="store_true", help="N>1: exchange first, then one local SpMM")
    p.add_argument("--transport", choices=["p2p", "nccl"], default="p2p",
                   help="N>1, lshape: direct stores into peer symmetric memory, or NCCL all_to_all of packed rows")
    p.add_argument("--push-ctas", type=int, default=0, help="CTAs of the p2p push kernel (0: library default)")
    p.add_argument("--hubs", type=int, default=0,
                   help="N>1, lshape: the lowest K ranks exchange through NVSwitch multicast (multimem broadcast of "
                        "their X block, in-switch reduction of their partial rows)")
    p.add_argument("--refine", action="store_true",
                   help="N>1, lshape: descend on the step timeline model from the cost-cap bounds (measured: no gain "
                        "at p = 8 on R-MAT 24, 7.5 vs 7.3 ms, so it is off by default)")
    p.add_argument("--comm-weight", type=float, default=0.0,
                   help="N>1, lshape: cost of one row moved, in nnz-equivalents, in the partition's balance")
    p.add_argument("--two-comms", action="store_true", help="N>1, lshape: partial rows on a second NCCL communicator")
    p.add_argument("--partition", choices=["lshape", "rows"], default="lshape",
                   help="N>1: L-shaped partition (blocks above the diagonal are multiplied at the column owner and "
                        "their partial rows sent back) or plain row blocks with a pull exchange")
    p.add_argument("--diag", action="store_true", help="N>1: print per-rank timings of the step's parts to stderr")
    p.add_argument("--row-weight", type=float, default=6.0,
                   help="partition balance: cost(row) = nnz(row) + row_weight; 6 = measured per-row overhead of the "
                        "SpMM kernel in edge-equivalents (DESIGN.md, multi-GPU)")
    p.add_argument("--cpu-iters", type=int, default=3, help="iterations of the CPU baseline sample")
    return p.parse_args()


# ------------------------------------------------------------------------------------------------------
# shared plumbing
# ------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.proc = None
        self.lines = []
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])), mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            m = json.load(f)
        return float(m["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def spmm_alg_bytes(nnz, n_rows, d, elem=4, mul=False, add=False):
    """SURVEY §8(d) gather model, no reuse credited: nnz*(col + val + d*x) + N*d*y + (N+1)*ptr
    (+ nnz*d for the fused Hadamard operand, + 2*N*d for the fused e1*e2 addend)."""
    b = nnz * (4 + elem + d * elem) + n_rows * d * elem + (n_rows + 1) * 4
    if mul:
        b += nnz * d * elem
    if add:
        b += 2 * n_rows * d * elem
    return b


def msi_inputs():
    """MSI-shaped graph + synthetic node2vec-like embeddings (see synth.py; the real blobs are not
    distributed with the reference: .MISSING_LARGE_BLOBS)."""
    import scipy.sparse as sp
    import importlib
    synth = importlib.import_module("gcn-drug-repurposing_b200.synth")
    g = synth.msi_synth()
    n = g["n"]
    a = sp.csr_matrix((g["weight"], (g["src"], g["dst"])), shape=(n, n))
    a.sort_indices()
    feats = synth.synth_embeddings(n, MSI_D, seed=1)
    return a, feats


def epoch_batches(n, batch, epochs, seed):
    """DataLoader(shuffle=True) equivalent: a fresh permutation per epoch, split into batches."""
    import torch
    gen = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(epochs):
        perm = torch.randperm(n, generator=gen)
        out.append([perm[i:i + batch] for i in range(0, n, batch)])
    return out


# ------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own library calls (oracle/gcn_oracle.py restates modules/model.py + train.py
# on torch CPU; /root/reference itself does not travel to the GPU box).  Used for cpu_baseline and for
# --impl reference.  This is the only place bench.py executes oracle/.
# ------------------------------------------------------------------------------------------------------
class CpuReferenceGCN:
    def __init__(self, a, feats):
        import torch
        from oracle import gcn_oracle
        self.torch, self.o = torch, gcn_oracle
        a_hat = gcn_oracle.preprocess_graph(a)
        self.adj = gcn_oracle.to_sparse_tensor(a_hat)
        self.nnz = a_hat.nnz
        np.random.seed(1)
        torch.manual_seed(1)
        self.tr = gcn_oracle.Trainer(feats, self.adj, feats.shape[1], num_layers=MSI_LAYERS, layer_decay=MSI_DECAY,
                                     lr=MSI_LR)
        self.n = feats.shape[0]
        self.beta = None

    def set_beta(self, beta=None):
        if beta is None:
            # train.py:165-167 on the first forward output.  The N×N fp32 Gram + np.percentile costs ~20 s and
            # 7 GB on the host; it is a one-off outside every timed region.
            h0 = self.tr.forward().detach()
            beta = float(self.o.beta_percentile(h0, MSI_BETA_PCT))
        self.beta = beta
        return beta

    def iteration(self, batch):
        return self.tr.step(batch, self.beta)[0]


def time_cpu_iterations(cpu, batches, iters):
    t0 = time.perf_counter()
    for b in batches[:iters]:
        cpu.iteration(b)
    return time.perf_counter() - t0


# ------------------------------------------------------------------------------------------------------
# our arm, workload msi_gcn
# ------------------------------------------------------------------------------------------------------
def run_msi_gcn(args):
    import importlib
    import torch
    gdr = importlib.import_module("gcn-drug-repurposing_b200")
    from gdr_b200 import trainer as trainer_mod
    from gdr_b200.graph import DeviceCSR

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    a, feats = msi_inputs()
    n, d = feats.shape
    a_hat = gdr.preprocess_graph(a)                                  # CUDA: D^-1/2 (A+I) D^-1/2
    adj = gdr.convert_sparse_matrix_to_sparse_tensor(a_hat)
    nnz = a_hat.nnz
    np.random.seed(1)
    torch.manual_seed(1)
    tr = trainer_mod.GCNTrainer(feats, adj, hidden_units=d, num_layers=MSI_LAYERS, layer_decay=MSI_DECAY,
                                lr=MSI_LR, batch_size=MSI_BATCH, beta_percentile=MSI_BETA_PCT, device="cuda:0")
    iters_per_epoch = (n + MSI_BATCH - 1) // MSI_BATCH
    total_epochs = args.warmup + args.steps
    host_batches = epoch_batches(n, MSI_BATCH, 2 * total_epochs + 2, seed=1)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)

    # ---- device-resident arm: features, adjacency and index batches already in HBM -------------------
    dev_batches = [[b.to(dev) for b in ep] for ep in host_batches[:total_epochs]]
    for ep in dev_batches[:args.warmup]:
        tr.epoch(ep)
    torch.cuda.synchronize()
    clocks = ClockSampler(0)
    clocks.start()
    DeviceCSR.timer = []
    gdr._ffi.launch_count_reset()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    torch.cuda.synchronize()
    wall0 = time.perf_counter()
    for (e0, e1), ep in zip(ev, dev_batches[args.warmup:]):
        flush.zero_()                                                # L2 flush between timed steps (untimed)
        e0.record()
        tr.epoch(ep)
        e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - wall0
    launches = gdr._ffi.launch_count()
    step_ms = [a_.elapsed_time(b_) for a_, b_ in ev]
    ms_per_step = float(np.sum(step_ms) / args.steps)
    spmm_events = DeviceCSR.timer
    DeviceCSR.timer = None
    spmm_ms = sum(e0.elapsed_time(e1) for e0, e1, *_ in spmm_events)
    spmm_bytes = sum(spmm_alg_bytes(nz, nr, dd, 4, mul, add) for _, _, nz, nr, dd, mul, add in spmm_events)
    spmm_edges = sum(nz for _, _, nz, *_ in spmm_events)

    # ---- end-to-end arm: host buffers through the public API, as train.py:157-160 moves them -----------
    feats_pinned = tr.dataset.features.pin_memory()
    e2e_batches = [[b.pin_memory() for b in ep] for ep in host_batches[total_epochs:2 * total_epochs]]
    h2d = d2h = 0

    def e2e_epoch(ep, count):
        nonlocal h2d, d2h
        loss_val = None
        for b in ep:
            bd = b.to(dev, non_blocking=True)
            xd = feats_pinned.to(dev, non_blocking=True)             # train.py:159 re-uploads the features
            loss = tr.iteration(bd, features=xd)
            loss_val = loss.item()                                   # D2H read of the step result
            if count:
                h2d += b.numel() * 8 + feats_pinned.numel() * 4
                d2h += 4
        return loss_val

    for ep in e2e_batches[:args.warmup]:
        e2e_epoch(ep, False)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for ep in e2e_batches[args.warmup:]:
        e2e_epoch(ep, True)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    clk = clocks.stop()

    peak, peak_src = measured_peaks()
    achieved = spmm_bytes / (spmm_ms * 1e-3) / 1e9 if spmm_ms > 0 else 0.0
    line = {
        "metric": "gcn_epochs_per_s", "value": 1000.0 / ms_per_step, "unit": "epochs/s", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "msi_gcn: 2-layer GSS GCN hidden 256 on the MSI-shaped weighted graph "
                               "(BASELINE configs[1]); step = 1 epoch = %d iterations of fwd+loss+bwd+Adam" % iters_per_epoch,
                   "nodes": n, "nnz_adj": int(nnz), "d": d, "batch": MSI_BATCH, "num_layers": MSI_LAYERS,
                   "iters_per_epoch": iters_per_epoch, "l2": "explicit 256 MiB flush between timed steps",
                   "graph": "MSI-synth (real blobs absent upstream: .MISSING_LARGE_BLOBS)"},
        "e2e": {"value": args.steps / e2e_s, "unit": "epochs/s", "h2d_bytes_per_step": h2d // args.steps,
                "d2h_bytes_per_step": d2h // args.steps},
        "gpu_launches": int(launches),
        "clocks": clk,
        "roofline": {"kernel": "spmm_warp_row_kernel<float> (all SpMM launches of the timed steps)", "bound": "hbm",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     # dram__bytes_read.sum + dram__bytes_write.sum per launch of this kernel at this size, from the
                     # ncu --set full capture profiles/msi_iteration_r1_full_raw.csv (39 MB read + 5..25 MB write)
                     "traffic": 54.0e6 if (feats.shape[0], feats.shape[1]) == (29959, 256) else None,
                     "peak_source": peak_src, "launches": len(spmm_events),
                     "avg_launch_ms": spmm_ms / max(len(spmm_events), 1),
                     "share_of_step": spmm_ms / (ms_per_step * args.steps),
                     "gedges_per_s": spmm_edges / (spmm_ms * 1e-3) / 1e9 if spmm_ms > 0 else 0.0,
                     "note": "X (N*d*4 = %.0f MB) is L2-resident at MSI scale: see spmm_rmat for the HBM-bound case"
                             % (n * d * 4 / 1e6)},
        "wall_s_timed": wall,
    }
    return line, (a, feats, tr.beta, host_batches)


# ------------------------------------------------------------------------------------------------------
# R-MAT SpMM (single GPU secondary measurement, and the N>1 workload)
# ------------------------------------------------------------------------------------------------------
def build_rmat_csr(scale, edge_factor, dev, seed=1, row_range=None):
    """Graph500 R-MAT edges -> coalesced (duplicates merged), (A+I)-normalised fp32 DeviceCSR on `dev`.
    row_range=(lo, hi) keeps only those rows (the rank's partition), columns stay global."""
    import importlib
    import torch
    gdr = importlib.import_module("gcn-drug-repurposing_b200")
    n = 1 << scale
    r, c = gdr.synth.rmat_coo(scale, edge_factor, seed, dev)
    keys = torch.unique(r * n + c)                    # sorted, duplicates merged (post-coalescing nnz)
    del r, c
    rows = torch.div(keys, n, rounding_mode="floor")
    cols = keys - rows * n
    del keys
    g = torch.Generator(device=dev).manual_seed(seed + 1)
    vals = 1.0 - torch.rand(rows.numel(), generator=g, device=dev, dtype=torch.float64)   # U(0,1]
    csr = gdr.DeviceCSR.from_coo(rows, cols, vals, n, n)
    del rows, cols, vals
    a_hat = csr.normalized_with_identity()            # fp64, helpers/helper.py:82-89 on the device
    del csr
    out = a_hat.astype(torch.float32)
    if row_range is not None:
        lo, hi = row_range
        p0, p1 = int(out.rowptr[lo].item()), int(out.rowptr[hi].item())
        out = gdr.DeviceCSR((out.rowptr[lo:hi + 1] - p0).contiguous(), out.col[p0:p1].contiguous(),
                            out.val[p0:p1].contiguous(), hi - lo, n)
    torch.cuda.empty_cache()
    return out


def run_rmat_spmm_single(args, steps, warmup):
    """Y = Â·X on one GPU; X (N*d*4 bytes) >> L2, so consecutive launches need no explicit flush."""
    import torch
    dev = torch.device("cuda", 0)
    scale = args.rmat_scale
    csr = None
    while csr is None:
        try:
            csr = build_rmat_csr(scale, args.rmat_edge_factor, dev)
        except torch.OutOfMemoryError:
            torch.cuda.empty_cache()
            scale -= 1
    n, d = csr.n_rows, args.rmat_d
    x = torch.randn(n, d, device=dev, generator=torch.Generator(device=dev).manual_seed(2))
    y = torch.empty_like(x)
    for _ in range(warmup):
        csr.spmm(x, out=y)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        csr.spmm(x, out=y)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    peak, peak_src = measured_peaks()
    b_alg = spmm_alg_bytes(csr.nnz, n, d)
    b_min = csr.nnz * 8 + 2 * n * d * 4 + (n + 1) * 4
    gbs = b_alg / (ms * 1e-3) / 1e9
    return {"workload": f"rmat scale {scale} edge-factor {args.rmat_edge_factor} d {d} fp32, (A+I)-normalised",
            "nodes": n, "nnz": csr.nnz, "ms_per_spmm": ms, "gedges_per_s": csr.nnz / (ms * 1e-3) / 1e9,
            "alg_gb": b_alg / 1e9, "compulsory_gb": b_min / 1e9, "achieved_gbs": gbs, "peak_gbs": peak,
            "frac": gbs / peak, "frac_compulsory": b_min / (ms * 1e-3) / 1e9 / peak, "peak_source": peak_src,
            "l2": "inputs larger than L2 (X = %.1f GB)" % (n * d * 4 / 1e9),
            # ncu --set full of the same kernel at scale 24, d 128 (profiles/spmm_rmat24_r1b_hub_split_full_raw.csv):
            # 166.9 GB read + 9.2 GB write per launch at 5.28 TB/s = 81 % of the measured HBM peak
            "traffic": 176.14e9 if (scale, d, args.rmat_edge_factor) == (24, 128, 32) else None}


def run_rmat_spmm_dist(args):
    """Row-partitioned SpMM: rank r owns rows [lo_r, hi_r) of Â (nnz-balanced) and the matching rows of X
    and Y.  One step = all-gather of the X blocks over NCCL/NVLink + the local SpMM (dist.py)."""
    import importlib
    import torch
    import torch.distributed as dist
    gdr = importlib.import_module("gcn-drug-repurposing_b200")
    from gdr_b200 import dist as gdist
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)        # exchanges overtake queued SpMM CTAs
    import datetime
    dist.init_process_group("nccl", device_id=dev, pg_options=opts, timeout=datetime.timedelta(seconds=180))
    full = build_rmat_csr(args.rmat_scale, args.rmat_edge_factor, dev)        # same seed on every rank
    n, d, nnz_total = full.n_rows, args.rmat_d, full.nnz
    if args.partition == "lshape":
        deg = (full.rowptr[1:] - full.rowptr[:-1]).to(torch.int64)
        row_of = torch.repeat_interleave(torch.arange(n, device=dev), deg)
        hist, bin_size = gdist.coarse_histogram(row_of, full.col.to(torch.int64), n, min(1024, n))
        del row_of, deg
        hist_np = hist.cpu().numpy()
        bounds = gdist.lshape_bounds(hist_np, bin_size, n, world, row_weight=args.row_weight,
                                     comm_weight=args.comm_weight)
        if args.refine:
            bounds, _pred = gdist.refine_lshape_bounds(hist_np, bin_size, n, bounds, d)
    else:
        rp = full.rowptr.cpu().numpy().astype(np.int64)
        bounds = gdist.nnz_balanced_bounds(rp + (np.arange(len(rp)) * args.row_weight).astype(np.int64), world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    gen = torch.Generator(device=dev).manual_seed(2)
    x_full = torch.randn(n, d, device=dev, generator=gen)                     # identical on every rank
    y_full = full.spmm(x_full)
    y_ref = y_full[lo:hi].clone()                                             # 1-GPU product, for the check below
    t1a, t1b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t1a.record()
    for _ in range(3):
        full.spmm(x_full, out=y_full)                                         # the same product on ONE GPU, timed
    t1b.record()
    torch.cuda.synchronize()
    ms_1gpu = t1a.elapsed_time(t1b) / 3
    del y_full
    if args.partition == "lshape":
        transport = args.transport
        if transport == "p2p" and not gdist.p2p_available(dev):
            transport = "nccl"                      # no peer mapping on this box: NCCL all-to-all, two communicators
            args.two_comms, args.hubs = True, 0
        g2 = dist.new_group(backend="nccl", pg_options=dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)) \
            if (args.two_comms and transport == "nccl") else None
        if args.hubs > 0:
            part = gdist.NvlsLShapedSpMM.from_full(full, bounds, rank, world, d, n_hubs=args.hubs,
                                                   push_ctas=args.push_ctas)
        else:
            part = gdist.LShapedSpMM.from_full(full, bounds, rank, world, d, group_partials=g2,
                                               transport=transport, push_ctas=args.push_ctas)
    else:
        part = gdist.RowPartitionedSpMM.from_full(full, bounds, rank, world, d, overlap=not args.no_overlap)
    del full
    part.x_own.copy_(x_full[lo:hi])
    del x_full
    torch.cuda.empty_cache()
    x_local = None
    y = part.spmm(x_local)
    err = (y - y_ref).abs()
    tol_ok = bool(((err <= 1e-6) | (err <= 1e-5 * y_ref.abs()) | (err <= 4e-7 * y_ref.abs().max())).all().item())
    max_err = float(err.max().item())
    del y_ref, err
    for _ in range(args.warmup):
        part.spmm(x_local)
    torch.cuda.synchronize()
    dist.barrier()
    if args.diag:
        def timed(fn, reps=5):
            torch.cuda.synchronize()
            dist.barrier()
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a_.record()
            for _ in range(reps):
                fn()
            b_.record()
            torch.cuda.synchronize()
            return a_.elapsed_time(b_) / reps
        if args.partition == "lshape":
            t_x = timed(part.exchange_x)
            nop = lambda: None                     # every rank must enter timed() (it holds a barrier)
            t_push = timed(part.push_product)
            t_p = timed(part.exchange_partials)
            t_pull = timed(lambda: part.pull.spmm(part.xbuf, out=part.y))
            t_add = timed((lambda: part.add.spmm(part.precv, e1=part.y, out=part.y)) if part.n_precv else nop)
            t_red = timed(part.hub_reduce)
            t_all = timed(lambda: part.spmm(None))
            sys.stderr.write("[diag] rank %d rows %d | x-exchange %.2f ms (recv %d send %d rows) | push %.2f ms (nnz %d, "
                             "%d rows) | partial-exchange %.2f ms (recv %d) | pull %.2f ms (nnz %d) | add %.2f ms | "
                             "hub-reduce %.2f ms | step %.2f ms\n" % (
                                 rank, hi - lo, t_x, sum(part.xrecv_splits), int(part.send_idx.numel()), t_push,
                                 part.push.nnz, part.n_psend, t_p, part.n_precv, t_pull, part.pull.nnz, t_add, t_red,
                                 t_all))
            sys.stderr.flush()
        t_ex = 0.0 if args.partition == "lshape" else timed(part._exchange)
        if args.partition == "lshape":
            extra = ""
        elif part.overlap:
            t_own = timed(lambda: part.csr_own.spmm(part.xbuf, out=part.y))
            t_rem = timed(lambda: part.csr_rem.spmm(part.xbuf, e1=part.y, out=part.y))
            extra = "own %.2f ms (nnz %d) rem %.2f ms (nnz %d)" % (t_own, part.csr_own.nnz, t_rem, part.csr_rem.nnz)
        else:
            t_all = timed(lambda: part.csr_all.spmm(part.xbuf, out=part.y))
            extra = "local spmm %.2f ms (nnz %d)" % (t_all, part.csr_all.nnz)
        if args.partition != "lshape":
            sys.stderr.write("[diag] rank %d rows %d recv_rows %d send_rows %d exchange %.2f ms %s\n" % (
                rank, hi - lo, part.recv_rows, int(part.send_idx.numel()), t_ex, extra))
            sys.stderr.flush()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    gdr._ffi.launch_count_reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        part.spmm(x_local)
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    launches = torch.tensor([gdr._ffi.launch_count()], device=dev, dtype=torch.int64)
    dist.all_reduce(launches)
    ms = float(ms.item())
    # ---- end-to-end arm: the rank's X block comes from pinned host memory and its Y block goes back to it,
    # every step, inside the timed region (host buffers through the public call) --------------------------------
    x_host = torch.empty(part.x_own.shape, dtype=torch.float32).pin_memory()
    x_host.copy_(part.x_own)
    y_host = torch.empty(part.y.shape, dtype=torch.float32).pin_memory()
    e2e_steps = max(2, min(args.steps, 4))

    def e2e_step():
        part.x_own.copy_(x_host, non_blocking=True)
        yb = part.spmm(None)
        y_host.copy_(yb, non_blocking=True)

    e2e_step()
    torch.cuda.synchronize()
    dist.barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(e2e_steps):
        e2e_step()
    f1.record()
    torch.cuda.synchronize()
    dist.barrier()
    e2e_ms = torch.tensor([f0.elapsed_time(f1) / e2e_steps], device=dev, dtype=torch.float64)
    dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    io = torch.tensor([x_host.numel() * 4, y_host.numel() * 4], device=dev, dtype=torch.int64)
    dist.all_reduce(io)
    e2e_ms = float(e2e_ms.item())
    line = None
    if rank == 0:
        clk = clocks.stop()
        peak, peak_src = measured_peaks()
        b_alg = spmm_alg_bytes(nnz_total, n, d)
        gbs = b_alg / (ms * 1e-3) / 1e9
        line = {
            "metric": "spmm_gedges_per_s", "value": nnz_total / (ms * 1e-3) / 1e9, "unit": "GEdges/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "rmat_spmm: Y = A_hat X, R-MAT scale %d edge-factor %d d %d, rows partitioned over the GPUs, "
                                   "exchange of the referenced X rows / partial Y rows + local SpMM per step "
                                   "(BASELINE configs[4])"
                                   % (args.rmat_scale, args.rmat_edge_factor, d),
                       "nodes": n, "nnz": nnz_total, "d": d,
                       "parallelism": f"{'L-shaped' if args.partition == 'lshape' else 'row'}-partition x{world}",
                       "same_product_on_1_gpu_ms": ms_1gpu, "bounds": [int(b) for b in bounds],
                       "l2": "inputs larger than L2", "transport": part.transport,
                       "rows_received_per_rank_rank0": part.recv_rows,
                       "check_vs_1gpu_product": {"within_tolerance": tol_ok, "max_abs_err": max_err}},
            "e2e": {"value": nnz_total / (e2e_ms * 1e-3) / 1e9, "unit": "GEdges/s", "ms_per_step": e2e_ms,
                    "steps": e2e_steps, "h2d_bytes_per_step": int(io[0].item()), "d2h_bytes_per_step": int(io[1].item()),
                    "note": "every rank uploads its X block from pinned host memory and downloads its Y block, each step"},
            "gpu_launches": int(launches.item()), "clocks": clk,
            "roofline": {"kernel": "spmm_warp_row_kernel<float> (push + pull + add products of every rank) with the exchange "
                                   "(push_rows_kernel / NCCL) inside the step", "bound": "hbm",
                         "achieved": gbs, "peak": peak * world, "unit": "GB/s", "frac": gbs / (peak * world),
                         "traffic": None, "peak_source": peak_src + f" x{world} GPUs",
                         "exchange_bytes_in_rank0": int(part.recv_rows * d * 4),
                         "allgather_bytes_in_would_be": int((n - (hi - lo)) * d * 4)},
        }
    dist.destroy_process_group()
    return line


# ------------------------------------------------------------------------------------------------------
# reference arm
# ------------------------------------------------------------------------------------------------------
def run_reference(args, workload):
    """The reference's own CPU implementation of the path (torch CPU sparse.mm / mm / Adam exactly as
    modules/model.py + train.py call them; restated in oracle/gcn_oracle.py because /root/reference does not
    travel to the GPU box), all host threads.  Each step is a bounded sample of the workload."""
    import torch
    # torchrun exports OMP_NUM_THREADS=1; the reference arm is the only work on the box while it runs (the other
    # ranks exit at once), so it gets every host thread
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    cores = torch.get_num_threads()
    if workload == "msi_gcn":
        a, feats = msi_inputs()
        n = feats.shape[0]
        iters_per_epoch = (n + MSI_BATCH - 1) // MSI_BATCH
        cpu = CpuReferenceGCN(a, feats)
        cpu.set_beta()
        flat = [b for ep in epoch_batches(n, MSI_BATCH, 1 + (args.warmup + args.steps) // iters_per_epoch + 1, 1)
                for b in ep]
        # a step of this arm = ONE iteration (1/15 epoch), scaled to epochs/s
        for b in flat[:args.warmup]:
            cpu.iteration(b)
        times = []
        for b in flat[args.warmup:args.warmup + args.steps]:
            t0 = time.perf_counter()
            cpu.iteration(b)
            times.append(time.perf_counter() - t0)
        s_per_iter = float(np.mean(times))
        value = 1.0 / (s_per_iter * iters_per_epoch)
        sample = "1 of the %d iterations of an epoch per step (full graph, batch %d), scaled to epochs/s" % (
            iters_per_epoch, MSI_BATCH)
        return {"impl": "reference", "metric": "gcn_epochs_per_s", "value": value, "unit": "epochs/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": s_per_iter * 1e3,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "msi_gcn (same inputs as the GPU arm)", "nodes": n, "nnz_adj": int(cpu.nnz),
                           "d": MSI_D, "batch": MSI_BATCH, "iters_per_epoch": iters_per_epoch},
                "cpu_baseline": {"value": value, "unit": "epochs/s", "cores": cores, "kind": "port", "sample": sample},
                "e2e": {"value": value, "unit": "epochs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
    # rmat_spmm: torch.sparse.mm on the host, on a smaller R-MAT of the same family (bounded sample)
    import importlib
    synth = importlib.import_module("gcn-drug-repurposing_b200.synth")
    from oracle import gcn_oracle
    import scipy.sparse as sp
    scale = min(args.rmat_scale, 20)
    n = 1 << scale
    r, c = synth.rmat_coo(scale, args.rmat_edge_factor, 1, "cpu")
    keys = torch.unique(r * n + c).numpy()
    m = sp.csr_matrix((np.random.default_rng(2).uniform(0.0, 1.0, len(keys)) + 1e-9, (keys // n, keys % n)),
                      shape=(n, n))
    adj = gcn_oracle.to_sparse_tensor(gcn_oracle.preprocess_graph(m))
    x = torch.randn(n, args.rmat_d)
    for _ in range(min(args.warmup, 2)):
        torch.sparse.mm(adj, x)
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        torch.sparse.mm(adj, x)
        times.append(time.perf_counter() - t0)
    value = adj._nnz() / float(np.mean(times)) / 1e9
    return {"impl": "reference", "metric": "spmm_gedges_per_s", "value": value, "unit": "GEdges/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(np.mean(times)) * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "rmat_spmm via torch.sparse.mm on the host", "nodes": n, "nnz": adj._nnz(),
                       "d": args.rmat_d},
            "cpu_baseline": {"value": value, "unit": "GEdges/s", "cores": cores, "kind": "port",
                             "sample": "R-MAT scale %d (same generator, edge-factor, d) instead of scale %d" % (
                                 scale, args.rmat_scale)},
            "e2e": {"value": value, "unit": "GEdges/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def main():
    args = parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    workload = args.workload
    if workload == "auto":
        workload = "msi_gcn" if max(world, args.gpus) == 1 else "rmat_spmm"
    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(run_reference(args, workload)), flush=True)
        return 0
    if workload == "msi_gcn":
        if rank != 0:
            return 0      # the MSI graph runs on one GPU (north_star); nothing to shard
        line, (a, feats, beta, host_batches) = run_msi_gcn(args)
        if not args.no_rmat:
            line["spmm_rmat"] = run_rmat_spmm_single(args, steps=10, warmup=3)
        if not args.no_cpu_baseline:
            import torch
            cpu = CpuReferenceGCN(a, feats)
            cpu.set_beta(float(beta))           # same beta as the GPU arm; the percentile is not part of the step
            flat = [b for ep in host_batches for b in ep]
            cpu.iteration(flat[0])              # warm-up
            secs = time_cpu_iterations(cpu, flat[1:], args.cpu_iters)
            ipe = line["config"]["iters_per_epoch"]
            line["cpu_baseline"] = {"value": args.cpu_iters / secs / ipe, "unit": "epochs/s",
                                    "cores": torch.get_num_threads(), "kind": "port",
                                    "sample": "%d of the %d iterations of one epoch (after 1 warm-up), %.1f s of CPU work, "
                                              "torch CPU restatement of modules/model.py + train.py (oracle/gcn_oracle.py)"
                                              % (args.cpu_iters, ipe, secs)}
        print(json.dumps(line), flush=True)
        return 0
    if world == 1:
        line = run_rmat_spmm_single(args, args.steps, args.warmup)
        peak, src = measured_peaks()
        out = {"metric": "spmm_gedges_per_s", "value": line["gedges_per_s"], "unit": "GEdges/s", "n_gpus": 1,
               "steps": args.steps, "warmup": args.warmup, "ms_per_step": line["ms_per_spmm"], "higher_is_better": True,
               "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": "rmat_spmm: " + line["workload"], "l2": line["l2"]}, "e2e": None,
               "gpu_launches": args.steps,
               "roofline": {"kernel": "spmm_warp_row_kernel<float>", "bound": "hbm", "achieved": line["achieved_gbs"],
                            "peak": peak, "unit": "GB/s", "frac": line["frac"], "traffic": None, "peak_source": src},
               "spmm_rmat": line}
        print(json.dumps(out), flush=True)
        return 0
    line = run_rmat_spmm_dist(args)
    if line is not None:
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
