#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 GCN-propagation path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload ...]

Workloads (BASELINE.json `metric`: "GCN epochs/s on MSI graph; SpMM GEdges/s + HBM GB/s at 1/2/4/8 B200"):

  msi_gcn    (default at N=1; BASELINE configs[1]; the same line carries `spmm_rmat` (RMAT-24 SpMM, the HBM-bound
             kernel number), `rwr` (configs[2]: all diffusion profiles) and `rmat20_gcn` (configs[3]))
             2-layer GSS GCN, hidden 256, on the MSI-shaped weighted
             graph (29,959 nodes / 478,728 undirected edges -> nnz(Â) = 987,415), batch 2048, Adam lr 3e-4,
             train.sh hyper-parameters.  One step = one epoch = ceil(N/B) = 15 iterations of
             (full-graph forward + batch GSS loss + backward + Adam), train.py:155-184.  metric = epochs/s.
             The same line carries `spmm_rmat`: the HBM-bound SpMM number on an R-MAT graph at 1 GPU.
  rmat_spmm  (default at N>1; BASELINE configs[4]) SpMM Y = Â·X on an R-MAT graph (scale 24, edge factor 32,
             d=128), L-shaped partition over the ranks (gcn-drug-repurposing_b200/dist.py): only the referenced
             X rows / the partial Y rows travel, as direct stores into peer symmetric memory over NVLink
             (--transport nccl: all_to_all of packed rows; --hub-first / --split-pull: reorder the step around
             rank 0's link; --partition rows: plain row blocks).  One step = one SpMM over the whole graph, exchanges included.
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
# multi-GPU defaults (measured choices)
DIST_ROW_WEIGHT = 6.0
DIST_HUB_FIRST = 0
DIST_SPLIT_PULL = 0
DIST_FUSED_PUSH = 0          # measured at N = 8: the incast on rank 0 stalls every sender's push product (10.0 vs 7.4 ms)
DIST_VBLOCKS_AT_8 = 16       # cyclic virtual-block layout at 8 GPUs: 6.69 ms vs 7.38 ms (profiles/sweep_r2d_n8_lshape_cyclic.jsonl)
DIST_REBALANCE = 0           # measured at N = 8 (profiles/sweep_r2b_n8_lshape_options.jsonl): ranks are coupled through the step's barriers, finish-time feedback does not help (7.8 vs 7.3 ms)

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
    p.add_argument("--hub-first", type=int, default=DIST_HUB_FIRST,
                   help="N>1, lshape/p2p: compute and send the partial rows for rank 0 before the rest of the push product")
    p.add_argument("--split-pull", type=int, default=DIST_SPLIT_PULL,
                   help="N>1, lshape/p2p: cut the pull product by source (rank 0's rows last, folded into the add pass), "
                        "per-source arrival signals instead of one barrier")
    p.add_argument("--no-rwr", action="store_true", help="skip the diffusion-profile measurement (configs[2]) at N=1")
    p.add_argument("--no-rmat20", action="store_true", help="skip the R-MAT-20 GCN measurement (configs[3]) at N=1")
    p.add_argument("--fused-push", type=int, default=DIST_FUSED_PUSH,
                   help="N>1, p2p: the push product stores its rows straight into the destination's receive region")
    p.add_argument("--rebalance", type=int, default=DIST_REBALANCE,
                   help="N>1, lshape: rounds of bounds refinement from the measured per-rank finish times (set-up)")
    p.add_argument("--vblocks", type=int, default=-1,
                   help="N>1, lshape: virtual blocks per rank, owned cyclically (0: plain L-shape; -1: 16 at 8 GPUs, else 0)")
    p.add_argument("--transposed", type=int, default=1, help="N>1: also measure A_hat^T G (second partition)")
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
    p.add_argument("--row-weight", type=float, default=DIST_ROW_WEIGHT,
                   help="partition balance: cost(row) = nnz(row) + row_weight, the per-row overhead of the SpMM kernel in "
                        "edge-equivalents (DESIGN.md, multi-GPU)")
    p.add_argument("--cpu-iters", type=int, default=24, help="iterations of the CPU baseline sample (median reported)")
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


def measure_l2_read_peak(dev, mib=32, reps=4):
    """L2 -> SM read bandwidth of this device in GB/s, measured now: every CTA of csrc/peaks.cu streams an L2-resident
    buffer (default 32 MiB, the size of the MSI feature matrix) with 128-bit loads; best of 5 launches, CUDA events."""
    import ctypes
    import importlib
    import torch
    gdr = importlib.import_module("gcn-drug-repurposing_b200")
    L = gdr._ffi.lib()
    fn = L.gdr_bench_l2_read                       # measurement utility: exported, not in include/gdr.h
    fn.restype = ctypes.c_int64
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    buf = torch.ones(mib << 20, dtype=torch.uint8, device=dev)
    sink = torch.zeros(64, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    best, moved = None, 0
    for i in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        moved = fn(buf.data_ptr(), buf.numel(), reps, 0, sink.data_ptr(), st)
        e1.record()
        torch.cuda.synchronize()
        if moved < 0:
            raise RuntimeError("gdr_bench_l2_read failed: %d" % moved)
        if i >= 2:                                  # two warm-up launches (first touch comes from HBM)
            t = e0.elapsed_time(e1)
            best = t if best is None else min(best, t)
    return {"gbs": moved / (best * 1e-3) / 1e9, "buffer_mib": mib, "bytes_per_launch": int(moved), "best_ms": best,
            "how": "csrc/peaks.cu l2_read_kernel: 4 CTAs/SM x 512 threads, LDG.E.128 non-allocating, every CTA reads the "
                   "whole L2-resident buffer %d times; best of 5 launches, measured in this run" % reps}


def measure_tf32_tflops(dev, n=8192, seconds=1.0):
    """Dense TF32 tensor-core throughput of this device, measured now the way MEASURED_PEAKS.json measures bf16:
    torch.matmul (cuBLAS) of two fp32 n x n matrices with TF32 allowed; burst = best single product, sustained = back to
    back for `seconds`."""
    import torch
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(n, n, device=dev)
        b = torch.randn(n, n, device=dev)
        c = torch.empty(n, n, device=dev)
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize()
        flops = 2.0 * n * n * n
        best = None
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b, out=c)
            e1.record()
            torch.cuda.synchronize()
            t = e0.elapsed_time(e1)
            best = t if best is None else min(best, t)
        k = max(10, int(seconds / (best * 1e-3)))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize()
        sustained = flops * k / (e0.elapsed_time(e1) * 1e-3) / 1e12
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    del a, b, c
    return {"burst_tflops": flops / (best * 1e-3) / 1e12, "sustained_tflops": sustained,
            "how": "torch.matmul fp32 %d^3 with allow_tf32 (cuBLAS), best of 10 / back to back for %.0f s, measured in "
                   "this run" % (n, seconds)}


_UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "sector": 1.0, "ns": 1e-9, "us": 1e-6,
         "ms": 1e-3, "s": 1.0, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}


def profile_traffic(csv_name, kernel_regex):
    """DRAM and L2 traffic per launch of a kernel from a committed `ncu --set full ... --page raw --csv` export
    under profiles/ (parsed at run time: the numbers in the JSON line cannot go stale without the file changing).
    Returns None if the file is missing or no launch matches."""
    import csv
    import re
    path = os.path.join(ROOT, "profiles", csv_name)
    if not os.path.exists(path):
        return None
    with open(path, newline="") as f:
        rows = list(csv.reader(f))
    if len(rows) < 3:
        return None
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    need = ["Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"]
    if any(k not in col for k in need):
        return None

    def val(r, k):
        try:
            return float(r[col[k]].replace(",", "")) * _UNIT.get(units[col[k]], 1.0)
        except (ValueError, IndexError):
            return float("nan")

    sel = [r for r in rows[2:] if len(r) > col["Kernel Name"] and re.search(kernel_regex, r[col["Kernel Name"]])]
    if not sel:
        return None
    out = {"file": "profiles/" + csv_name, "kernel": re.sub(r"\(.*", "", sel[0][col["Kernel Name"]]).replace("void ", ""),
           "launches_in_capture": len(sel),
           "file_mtime": time.strftime("%Y-%m-%d", time.gmtime(os.path.getmtime(path))),
           "dram_bytes_per_launch": float(np.mean([val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
                                                   for r in sel])),
           "duration_s_under_ncu": float(np.mean([val(r, "gpu__time_duration.sum") for r in sel]))}
    if "lts__t_sectors.sum" in col:
        out["l2_bytes_per_launch"] = float(np.mean([val(r, "lts__t_sectors.sum") * 32.0 for r in sel]))
    for k, name in (("lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts_throughput_pct"),
                    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_active_pct"),
                    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_throughput_pct")):
        if k in col:
            out[name] = float(np.mean([val(r, k) for r in sel]))
    return out


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


def msi_config(n, nnz, d, iters_per_epoch):
    """The workload description both arms print (identical dicts: the driver compares them)."""
    return {"workload": "msi_gcn: 2-layer GSS GCN hidden 256 on the MSI-shaped weighted graph (BASELINE configs[1]); "
                        "step = 1 epoch = %d iterations of fwd+loss+bwd+Adam" % iters_per_epoch,
            "nodes": int(n), "nnz_adj": int(nnz), "d": int(d), "batch": MSI_BATCH, "num_layers": MSI_LAYERS,
            "iters_per_epoch": int(iters_per_epoch),
            "step_of_each_arm": "GPU arm: one timed step = one full epoch (%d iterations); CPU reference arm: one timed "
                                "step = ONE iteration (1/%d epoch), its epochs/s = 1 / (%d x mean iteration time)"
                                % (iters_per_epoch, iters_per_epoch, iters_per_epoch),
            "l2": "GPU arm: explicit 256 MiB L2 flush between timed steps; CPU arm: the working set (features, "
                  "activations, adjacency: > 150 MB) exceeds the host caches",
            "graph": "MSI-synth (real blobs absent upstream: .MISSING_LARGE_BLOBS)"}


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
    from gdr_b200 import model as model_mod
    dev_batches = [[b.to(dev) for b in ep] for ep in host_batches[:total_epochs]]

    def timed_epochs(epochs):
        """(ms per epoch, SpMM events, libgdr launches) over `epochs`, L2 flushed before each (untimed)."""
        DeviceCSR.timer = []
        gdr._ffi.launch_count_reset()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in epochs]
        torch.cuda.synchronize()
        for (e0, e1), ep in zip(evs, epochs):
            flush.zero_()                                            # L2 flush between timed steps (untimed)
            e0.record()
            tr.epoch(ep)
            e1.record()
        torch.cuda.synchronize()
        events, DeviceCSR.timer = DeviceCSR.timer, None
        return float(np.mean([a_.elapsed_time(b_) for a_, b_ in evs])), events, gdr._ffi.launch_count()

    # `value` is measured with the layer-1 aggregations recomputed every iteration: that is the work the reference arm
    # does (train.py:159 uploads the features anew each iteration, so nothing could be reused there)
    hoist_available = bool(model_mod.HOIST_LAYER1)
    tr.hoist_layer1 = False
    for ep in dev_batches[:args.warmup]:
        tr.epoch(ep)
    torch.cuda.synchronize()
    clocks = ClockSampler(0)
    clocks.start()
    wall0 = time.perf_counter()
    ms_per_step, spmm_events, launches = timed_epochs(dev_batches[args.warmup:])
    wall = time.perf_counter() - wall0
    spmm_ms = sum(e0.elapsed_time(e1) for e0, e1, *_ in spmm_events)
    spmm_bytes = sum(spmm_alg_bytes(nz, nr, dd, 4, mul, add) for _, _, nz, nr, dd, mul, add in spmm_events)
    spmm_edges = sum(nz for _, _, nz, *_ in spmm_events)
    # extra, labelled: the same epochs with P1 = A_hat X and T1 = A_hat (P1*X) reused across iterations (they do not
    # depend on the weights; GCNTrainer's opt-in for the resident feature matrix it owns)
    ms_hoist = None
    if hoist_available:
        tr.hoist_layer1 = True
        tr.epoch(dev_batches[0])
        ms_hoist, _, _ = timed_epochs(dev_batches[args.warmup:args.warmup + max(2, min(args.steps, 5))])
        tr.hoist_layer1 = False

    # ---- end-to-end arm: host buffers through the public API, as train.py:157-160 moves them -----------
    feats_pinned = tr.dataset.features.pin_memory()
    e2e_batches = [[b.pin_memory() for b in ep] for ep in host_batches[total_epochs:2 * total_epochs]]
    h2d = d2h = 0

    def e2e_epoch(ep, count):
        """Host buffers through the public call (trainer.GCNTrainer.epoch_from_host): per iteration the index batch and
        the feature matrix are uploaded from pinned memory (train.py:157-160 moves both every iteration) and the loss
        is read back; the upload of the next iteration overlaps the current one on a copy stream."""
        nonlocal h2d, d2h
        losses = tr.epoch_from_host(ep, feats_pinned)
        if count:
            h2d += sum(b.numel() * 8 for b in ep) + len(ep) * feats_pinned.numel() * 4
            d2h += 4 * len(ep)
        return losses[-1] if losses else None

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
    prof = profile_traffic("msi_iteration_r3_full_raw.csv", r"spmm_warp_row_kernel<float, 2") or \
        profile_traffic("msi_iteration_r2_full_raw.csv", r"spmm_warp_row_kernel<float, 2") or \
        profile_traffic("msi_iteration_r1_full_raw.csv", r"spmm_warp_row_kernel<float, 2")
    same_shape = (feats.shape[0], feats.shape[1]) == (29959, 256)
    l2_meas = measure_l2_read_peak(dev)
    l2_peak = l2_meas["gbs"]
    n_spmm = max(len(spmm_events), 1)
    avg_launch_s = spmm_ms * 1e-3 / n_spmm
    traffic = (prof or {}).get("dram_bytes_per_launch") if same_shape else None
    # frac = DRAM bytes actually moved (ncu, per launch) over the live launch time, against the measured HBM peak; when no
    # capture matches this shape, the compulsory bytes (every operand once) stand in for the traffic
    compulsory = sum(nz * 8 + (2 + (1 if mul else 0) + (1 if add else 0)) * nr * dd * 4 + (nr + 1) * 4
                     for _, _, nz, nr, dd, mul, add in spmm_events) / n_spmm
    dram_gbs = (traffic if traffic else compulsory) / avg_launch_s / 1e9 if spmm_ms > 0 else 0.0
    l2_bytes = (prof or {}).get("l2_bytes_per_launch") if same_shape else None
    line = {
        "metric": "gcn_epochs_per_s", "value": 1000.0 / ms_per_step, "unit": "epochs/s", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": msi_config(n, nnz, d, iters_per_epoch),
        "layer1_hoist": {"value_uses_hoist": False,
                         "note": "value and e2e recompute the first layer's aggregations every iteration, like the "
                                 "reference arm; with_hoist reuses P1 = A_hat X and T1 = A_hat (P1*X), which do not "
                                 "depend on the weights, while the trainer's resident feature tensor is unchanged "
                                 "(GCNTrainer(hoist_layer1=True), the trainer's default)",
                         "epochs_per_s_with_hoist": (1000.0 / ms_hoist) if ms_hoist else None,
                         "ms_per_step_with_hoist": ms_hoist},
        "e2e": {"value": args.steps / e2e_s, "unit": "epochs/s", "h2d_bytes_per_step": h2d // args.steps,
                "d2h_bytes_per_step": d2h // args.steps},
        "gpu_launches": int(launches),
        "clocks": clk,
        "roofline": {"kernel": "spmm_warp_row_kernel<float> (all SpMM launches of the timed steps)", "bound": "hbm",
                     "achieved": dram_gbs, "peak": peak, "unit": "GB/s", "frac": dram_gbs / peak,
                     "achieved_basis": ("dram__bytes_read.sum + dram__bytes_write.sum per launch (ncu export named in "
                                        "traffic_source) / the launch time measured in this run" if traffic else
                                        "compulsory bytes (every operand once) / the launch time measured in this run"),
                     # SURVEY 8(d) gather model (every gathered X row counted as memory traffic): NOT a fraction of
                     # anything at this size, X sits in L2 and is re-read from there
                     "achieved_gather_gbs": achieved, "gather_model_bytes_per_launch": spmm_bytes / n_spmm,
                     "traffic": traffic,
                     "traffic_source": prof if same_shape else None,
                     "peak_source": peak_src, "launches": len(spmm_events),
                     "avg_launch_ms": spmm_ms / n_spmm,
                     "share_of_step": spmm_ms / (ms_per_step * args.steps),
                     "gedges_per_s": spmm_edges / (spmm_ms * 1e-3) / 1e9 if spmm_ms > 0 else 0.0,
                     "note": "X (N*d*4 = %.0f MB) is L2-resident at MSI scale: HBM is not the binding resource of this "
                             "kernel here (frac is small by construction), roofline_l2 is the bound that applies; "
                             "spmm_rmat is the HBM-bound case" % (n * d * 4 / 1e6)},
        # the bound that applies at MSI scale: gathered bytes come out of L2
        "roofline_l2": {"kernel": "spmm_warp_row_kernel<float> at MSI scale", "bound": "l2",
                        "achieved": achieved, "peak": l2_peak, "unit": "GB/s", "frac": achieved / l2_peak,
                        "achieved_basis": "gather-model bytes (what the SMs request from L2) / launch time of this run",
                        "peak_source": l2_meas,
                        "frac_by_ncu_l2_sectors": (l2_bytes / avg_launch_s / 1e9 / l2_peak) if l2_bytes else None,
                        "l2_bytes_per_launch_ncu": l2_bytes,
                        "lts_throughput_pct_ncu": (prof or {}).get("lts_throughput_pct") if same_shape else None},
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
    match = (scale, d, args.rmat_edge_factor) == (24, 128, 32)
    prof = profile_traffic("spmm_rmat24_r3_full_raw.csv", r"spmm_warp_row_kernel<float, 1") or \
        profile_traffic("spmm_rmat24_r2_full_raw.csv", r"spmm_warp_row_kernel<float, 1") or \
        profile_traffic("spmm_rmat24_r1b_hub_split_full_raw.csv", r"spmm_warp_row_kernel<float, 1")
    return {"workload": f"rmat scale {scale} edge-factor {args.rmat_edge_factor} d {d} fp32, (A+I)-normalised",
            "nodes": n, "nnz": csr.nnz, "ms_per_spmm": ms, "gedges_per_s": csr.nnz / (ms * 1e-3) / 1e9,
            "alg_gb": b_alg / 1e9, "compulsory_gb": b_min / 1e9, "achieved_gather_gbs": gbs, "peak_gbs": peak,
            # frac: DRAM bytes actually moved (ncu) over the live launch time / measured HBM peak; the SURVEY 8(d) gather
            # model credits no L2 reuse of hub columns and is reported as a rate, not as a fraction
            "frac": ((prof["dram_bytes_per_launch"] / (ms * 1e-3) / 1e9 / peak) if (prof and match) else None),
            "frac_compulsory": b_min / (ms * 1e-3) / 1e9 / peak, "peak_source": peak_src,
            "l2": "inputs larger than L2 (X = %.1f GB)" % (n * d * 4 / 1e9),
            # ncu --set full of the same kernel at scale 24, d 128, parsed from the committed export
            "traffic": (prof or {}).get("dram_bytes_per_launch") if match else None,
            "traffic_source": prof if match else None,
            "frac_dram_traffic": ((prof["dram_bytes_per_launch"] / (ms * 1e-3) / 1e9 / peak) if (prof and match) else None)}


def run_rmat_spmm_dist(args):
    """Partitioned SpMM: rank r owns rows [lo_r, hi_r) of Â and the matching rows of X and Y.  One step = exchange
    of the referenced X rows / partial Y rows over NVLink (peer stores or NCCL all-to-all) + the local products
    (dist.py, L-shaped partition by default)."""
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
    gen = torch.Generator(device=dev).manual_seed(2)
    x_full = torch.randn(n, d, device=dev, generator=gen)                     # identical on every rank
    transport = args.transport
    g2 = None
    if args.partition == "lshape":
        if transport == "p2p" and not gdist.p2p_available(dev):
            transport = "nccl"                      # no peer mapping on this box: NCCL all-to-all, two communicators
            args.two_comms = True
        g2 = dist.new_group(backend="nccl", pg_options=dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)) \
            if (args.two_comms and transport == "nccl") else None

    def finish_ms(part, reps=3):
        """Per-rank time from a common start until this rank's Y block is complete (waits on other ranks included)."""
        for _ in range(2):
            part.spmm(None)
        ts = []
        for _ in range(reps):
            torch.cuda.synchronize()
            dist.barrier()
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a_.record()
            part.spmm(None)
            b_.record()
            torch.cuda.synchronize()
            ts.append(a_.elapsed_time(b_))
        t = torch.tensor([float(np.median(ts))], device=dev, dtype=torch.float64)
        out = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return np.array([float(o.item()) for o in out])

    vblocks = args.vblocks if args.vblocks >= 0 else (DIST_VBLOCKS_AT_8 if world >= 8 else 0)
    if transport != "p2p" and args.vblocks < 0:
        vblocks = 0                                  # measured on the peer-store transport only

    def partition_and_check(mat, tag):
        """Bounds (+ measured rebalancing), the rank's partition object, the same product on ONE GPU (timed) and the
        check of the partitioned product against it on every rank."""
        hist_np = bin_size = None
        rebal = {"iterations": 0}
        if args.partition == "lshape":
            deg = (mat.rowptr[1:] - mat.rowptr[:-1]).to(torch.int64)
            row_of = torch.repeat_interleave(torch.arange(n, device=dev), deg)
            hist, bin_size = gdist.coarse_histogram(row_of, mat.col.to(torch.int64), n, min(1024, n))
            del row_of, deg
            hist_np = hist.cpu().numpy()
            del hist
        y_full = mat.spmm(x_full)
        t1a, t1b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t1a.record()
        for _ in range(3):
            mat.spmm(x_full, out=y_full)                                      # the same product on ONE GPU, timed
        t1b.record()
        torch.cuda.synchronize()
        ms_one = t1a.elapsed_time(t1b) / 3

        def build(bounds):
            if args.partition == "lshape":
                pt = gdist.LShapedSpMM.from_full(mat, bounds, rank, world, d, group_partials=g2, transport=transport,
                                                 push_ctas=args.push_ctas, hub_first=bool(args.hub_first),
                                                 split_pull=bool(args.split_pull), fused_push=bool(args.fused_push))
            else:
                pt = gdist.RowPartitionedSpMM.from_full(mat, bounds, rank, world, d, overlap=not args.no_overlap)
            pt.x_own.copy_(x_full[int(bounds[rank]):int(bounds[rank + 1])])
            return pt

        rows_sel = None
        if args.partition == "lshape" and vblocks > 0:
            # cyclic ownership of world * vblocks virtual L-shapes (hub blocks spread over all NVLink ports)
            vbounds = gdist.lshape_bounds(hist_np, bin_size, n, world * vblocks, row_weight=args.row_weight,
                                          comm_weight=args.comm_weight)
            part = gdist.LShapedSpMM.from_full_cyclic(mat, vbounds, rank, world, d, group_partials=g2, transport=transport,
                                                      push_ctas=args.push_ctas or 64, fused_push=bool(args.fused_push))
            rows_sel = part.rows_old
            part.x_own.copy_(x_full[rows_sel])
            bounds = [int(b) for b in part.bounds]
            rebal = {"iterations": 0, "virtual_blocks_per_rank": vblocks,
                     "virtual_block_bounds_natural_order": [int(b) for b in vbounds]}
        elif args.partition == "lshape":
            weights = np.ones(world)
            history = []
            for it in range(max(args.rebalance, 0) + 1):
                bounds = gdist.lshape_bounds(hist_np, bin_size, n, world, row_weight=args.row_weight,
                                             comm_weight=args.comm_weight, rank_weights=weights)
                if args.refine and it == 0:
                    bounds, _pred = gdist.refine_lshape_bounds(hist_np, bin_size, n, bounds, d)
                part = build(bounds)
                if it == max(args.rebalance, 0):
                    break
                t = finish_ms(part)                                           # identical array on every rank
                history.append({"bounds": [int(b) for b in bounds], "finish_ms_per_rank": [round(float(v), 3) for v in t]})
                weights = gdist.rebalance_weights(weights, t)
                del part
                torch.cuda.empty_cache()
            rebal = {"iterations": len(history), "history": history, "rank_weights": [round(float(w), 4) for w in weights],
                     "note": "bounds re-derived from the MEASURED per-rank finish time of a step (set-up, untimed)"}
        else:
            rp = mat.rowptr.cpu().numpy().astype(np.int64)
            bounds = gdist.nnz_balanced_bounds(rp + (np.arange(len(rp)) * args.row_weight).astype(np.int64), world)
            part = build(bounds)
        lo, hi = int(bounds[rank]), int(bounds[rank + 1])
        y_ref = (y_full[lo:hi] if rows_sel is None else y_full[rows_sel]).clone()
        del y_full
        torch.cuda.empty_cache()
        # check against the 1-GPU product on EVERY rank (reduced below): the partitioned product differs from it only
        # in fp32 summation order.  1st: the north_star tolerance (1e-5 rel / 1e-6 abs), as a fraction of the entries;
        # 2nd, explicit and relaxed: fp32 sums of up to ~10^5 terms, bounded relative to the row block's largest entry.
        y = part.spmm(None)
        err = (y - y_ref).abs()
        strict = (err <= 1e-6) | (err <= 1e-5 * y_ref.abs())
        relaxed = strict | (err <= 4e-7 * y_ref.abs().max())
        chk = torch.tensor([float(strict.double().mean().item()), float(relaxed.all().item()), -float(err.max().item())],
                           device=dev, dtype=torch.float64)
        dist.all_reduce(chk, op=dist.ReduceOp.MIN)                # worst rank
        # a second step on CHANGED inputs: the exchange buffers must not leak rows of the previous step
        part.x_own.mul_(-0.5)
        y2 = part.spmm(None)
        err2 = (y2 - (-0.5) * y_ref).abs()
        ok2 = torch.tensor([float(((err2 <= 1e-6) | (err2 <= 1e-5 * y_ref.abs()) | (err2 <= 4e-7 * y_ref.abs().max()))
                                  .all().item())], device=dev, dtype=torch.float64)
        dist.all_reduce(ok2, op=dist.ReduceOp.MIN)
        part.x_own.mul_(-2.0)
        res = {"bounds": [int(b) for b in bounds], "ms_1gpu": ms_one, "strict_frac": float(chk[0].item()),
               "tol_ok": bool(chk[1].item() > 0.5 and ok2.item() > 0.5), "max_err": -float(chk[2].item()),
               "rebalance": rebal}
        del y_ref, err, err2, y2, strict, relaxed
        return part, res

    def timed_steps(part):
        for _ in range(args.warmup):
            part.spmm(None)
        torch.cuda.synchronize()
        dist.barrier()
        a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a_.record()
        for _ in range(args.steps):
            part.spmm(None)
        b_.record()
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([a_.elapsed_time(b_) / args.steps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- the transposed product first (its objects are freed before the headline measurement) ---------------------
    transposed = None
    if args.transposed:
        full_t = full.transpose()                                             # explicit second CSR
        full._t = None
        full_t._t = None
        part_t, res_t = partition_and_check(full_t, "transposed")
        ms_t = timed_steps(part_t)
        transposed = {"workload": "G_in = A_hat^T G (the backward aggregation, SURVEY 8e): a second L-shaped partition of the "
                                  "explicit transposed CSR, same kernels and exchange",
                      "ms_per_step": ms_t, "gedges_per_s": nnz_total / (ms_t * 1e-3) / 1e9,
                      "same_product_on_1_gpu_ms": res_t["ms_1gpu"], "efficiency_vs_1gpu": res_t["ms_1gpu"] / (world * ms_t),
                      "bounds": res_t["bounds"], "rebalance": res_t["rebalance"],
                      "check_vs_1gpu_product": {"within_tolerance": res_t["tol_ok"], "max_abs_err": res_t["max_err"],
                                                "strict_1e-5rel_1e-6abs_fraction_worst_rank": res_t["strict_frac"],
                                                "ranks_checked": world}}
        tol_ok_t = res_t["tol_ok"]
        del part_t, full_t
        torch.cuda.empty_cache()
    else:
        tol_ok_t = True
    part, res = partition_and_check(full, "forward")
    bounds, ms_1gpu = res["bounds"], res["ms_1gpu"]
    strict_frac, tol_ok, max_err = res["strict_frac"], res["tol_ok"] and tol_ok_t, res["max_err"]
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    del full, x_full
    torch.cuda.empty_cache()
    x_local = None
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

            def fused_push_local():                # the fused push product with LOCAL outputs (compute share only)
                for lo_, hi_, _v in part._push_ranges:
                    part._push_product_rows(lo_, hi_, out=part._diag_psend[lo_:hi_])
            if getattr(part, "fused_push", False):
                part._diag_psend = torch.empty(part.n_psend, d, device=dev)
                t_push = timed(fused_push_local)
                t_p = 0.0
                timed(nop)
            else:
                t_push = timed(part.push_product)
                t_p = timed(part.exchange_partials)
            t_pull = timed(lambda: part.pull.spmm(part.xbuf, out=part.y))
            t_add = timed((lambda: part.add.spmm(part.precv, e1=part.y, out=part.y)) if part.n_precv else nop)
            t_all = timed(lambda: part.spmm(None))
            sys.stderr.write("[diag] rank %d rows %d | x-exchange %.2f ms (recv %d send %d rows) | push %.2f ms (nnz %d, "
                             "%d rows) | partial-exchange %.2f ms (recv %d) | pull %.2f ms (nnz %d) | add %.2f ms | "
                             "step %.2f ms\n" % (
                                 rank, hi - lo, t_x, sum(part.xrecv_splits), int(part.send_idx.numel()), t_push,
                                 part.push.nnz, part.n_psend, t_p, part.n_precv, t_pull, part.pull.nnz, t_add, t_all))
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
                       "parallelism": f"{'L-shaped' if args.partition == 'lshape' else 'row'}-partition x{world}" + (f", {vblocks} virtual blocks per rank owned cyclically" if vblocks else ""),
                       "same_product_on_1_gpu_ms": ms_1gpu,
                       "efficiency_vs_1gpu": ms_1gpu / (world * ms),          # E(p) = T(1) / (p T(p)), same process
                       "bounds": [int(b) for b in bounds], "row_weight": args.row_weight,
                       "fused_push": bool(getattr(part, "fused_push", False)), "rebalance": res["rebalance"],
                       "cyclic_virtual_blocks_per_rank": vblocks,
                       "hub_first": bool(getattr(part, "hub_first", False)),
                       "split_pull": bool(getattr(part, "_split_requested", False)),
                       "l2": "inputs larger than L2", "transport": part.transport,
                       "rows_received_per_rank_rank0": part.recv_rows,
                       "check_vs_1gpu_product": {"within_tolerance": tol_ok, "max_abs_err": max_err,
                                                 "strict_1e-5rel_1e-6abs_fraction_worst_rank": strict_frac,
                                                 "ranks_checked": world,
                                                 "second_step_on_changed_x": "checked (x -> -0.5 x)"}},
            "e2e": {"value": nnz_total / (e2e_ms * 1e-3) / 1e9, "unit": "GEdges/s", "ms_per_step": e2e_ms,
                    "steps": e2e_steps, "h2d_bytes_per_step": int(io[0].item()), "d2h_bytes_per_step": int(io[1].item()),
                    "note": "every rank uploads its X block from pinned host memory and downloads its Y block, each step"},
            "gpu_launches": int(launches.item()), "clocks": clk, "transposed": transposed,
            "roofline": {"kernel": "spmm_warp_row_kernel<float> (push + pull + add products of every rank) with the exchange "
                                   "(push_rows_kernel / NCCL) inside the step", "bound": "hbm",
                         # no multi-GPU ncu capture: achieved / frac count the compulsory bytes of the product (every
                         # operand once); the SURVEY 8(d) gather model is given as a rate, not as a fraction
                         "achieved": (nnz_total * 8 + 2 * n * d * 4 + (n + 1) * 4) / (ms * 1e-3) / 1e9,
                         "peak": peak * world, "unit": "GB/s",
                         "frac": (nnz_total * 8 + 2 * n * d * 4 + (n + 1) * 4) / (ms * 1e-3) / 1e9 / (peak * world),
                         "achieved_basis": "compulsory bytes / step time", "achieved_gather_gbs": gbs,
                         "traffic": None, "peak_source": peak_src + f" x{world} GPUs",
                         "exchange_bytes_in_rank0": int(part.recv_rows * d * 4),
                         "allgather_bytes_in_would_be": int((n - (hi - lo)) * d * 4)},
        }
    dist.destroy_process_group()
    if not tol_ok:
        if rank == 0 and line is not None:
            print(json.dumps(line), flush=True)
        sys.stderr.write("bench.py: the partitioned product is OUT OF TOLERANCE against the 1-GPU product\n")
        sys.exit(3)
    return line


# ------------------------------------------------------------------------------------------------------
# BASELINE configs[2]: diffusion profiles of ALL drugs and indications (one batched device solve)
# ------------------------------------------------------------------------------------------------------
class _SynthMSI:
    """Duck-typed MSI (multiscale/msi/msi.py:148-205) over the synthetic interactome arrays."""

    class _Graph:
        def __init__(self, g):
            self._g = g

        def edges(self, data=None, default=None):
            nm = self._g["names"]
            return ((nm[u], nm[v], float(w)) for u, v, w in zip(self._g["src"], self._g["dst"], self._g["weight"]))

        def __len__(self):
            return self._g["n"]

    def __init__(self, g, synth):
        nm = g["names"]
        self.graph = self._Graph(g)
        self.nodelist = list(nm)
        self.node2idx = {x: i for i, x in enumerate(nm)}
        t = g["node_type"]
        self.drugs_in_graph = [nm[i] for i in g["seeds"] if t[i] == synth.DRUG]
        self.indications_in_graph = [nm[i] for i in g["seeds"] if t[i] == synth.INDICATION]
        self.drug_or_indication2proteins = {nm[s]: set(nm[q] for q in g["seed_proteins"].get(int(s), []))
                                            for s in g["seeds"]}


def _rwr_cpu_worker(job):
    """One forked/spawned worker of the CPU arm: the reference algorithm (oracle/rwr_oracle.py) on a few seeds."""
    seed_ids, = job
    import importlib
    from oracle import rwr_oracle
    synth = importlib.import_module("gcn-drug-repurposing_b200.synth")
    g = _rwr_cpu_worker.g if hasattr(_rwr_cpu_worker, "g") else synth.msi_synth()
    _rwr_cpu_worker.g = g
    m0 = rwr_oracle.initial_matrix(g["n"], g["src"], g["dst"], g["weight"])
    sinks = [int(s) for s in g["seeds"]]
    t0 = time.perf_counter()
    for s in seed_ids:
        rwr_oracle.diffusion_profile(m0, int(s), g["seed_proteins"], sinks, synth.RWR_ALPHA, synth.RWR_TOL,
                                     synth.RWR_MAX_ITER)
    return time.perf_counter() - t0


def run_rwr(args, with_cpu=True):
    """All 2,501 seeds of the MSI-shaped interactome (29,959 nodes / 957,456 directed edges) in one batched fp64
    power iteration (csrc/rwr.cu), profiles left on the device; then the same with the D2H copy of the profiles;
    the reference algorithm on the host: 1 core, and cpu_count()/2 workers as predict_drug.py:93 forks them."""
    import importlib
    import torch
    gdr = importlib.import_module("gcn-drug-repurposing_b200")
    from oracle import rwr_oracle
    synth = gdr.synth
    g = synth.msi_synth()
    msi = _SynthMSI(g, synth)
    dp = gdr.DiffusionProfiles(alpha=synth.RWR_ALPHA, max_iter=synth.RWR_MAX_ITER, tol=synth.RWR_TOL, weights=None,
                               num_cores=4, save_load_file_path=None)
    dp.get_initial_M(msi)
    prob = dp._get_problem(msi)
    seeds = [msi.node2idx[s] for s in msi.drugs_in_graph + msi.indications_in_graph]
    n, nnz, ns = g["n"], int(dp.initial_M.nnz), len(seeds)
    prob.solve(seeds[:64], dp.alpha, dp.tol, dp.max_iter)                     # warm-up
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    x, iters = prob.solve(seeds, dp.alpha, dp.tol, dp.max_iter)
    e1.record()
    torch.cuda.synchronize()
    t_dev = e0.elapsed_time(e1) * 1e-3
    it = iters.cpu().numpy()
    sweeps = int(it.max())
    out_host = torch.empty(x.shape, dtype=x.dtype).pin_memory()
    t0 = time.perf_counter()
    x2, _ = prob.solve(seeds, dp.alpha, dp.tol, dp.max_iter)
    out_host.copy_(x2, non_blocking=True)
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    # SURVEY §8(d) bytes per sweep, S seeds batched: nnz*(4+8) + nnz*S*8 + 2*N*S*8 (fp64 gather model)
    b_sweep = nnz * 12 + nnz * ns * 8 + 2 * n * ns * 8
    peak, peak_src = measured_peaks()
    gbs = b_sweep * sweeps / t_dev / 1e9
    # check against the oracle on a few seeds (bit-for-bit on nearly every entry, same stopping iteration)
    m0 = rwr_oracle.initial_matrix(n, g["src"], g["dst"], g["weight"])
    sinks = [int(s) for s in g["seeds"]]
    worst, it_same = 0.0, True
    t0 = time.perf_counter()
    n_chk = 6
    for k in list(range(0, ns, max(ns // n_chk, 1)))[:n_chk]:
        ox, oit = rwr_oracle.diffusion_profile(m0, seeds[k], g["seed_proteins"], sinks, dp.alpha, dp.tol, dp.max_iter)
        worst = max(worst, float(np.abs(x[k].cpu().numpy() - ox).max() / np.abs(ox).max()))
        it_same = it_same and (oit == int(it[k]))
    t_cpu1 = (time.perf_counter() - t0) / n_chk
    out = {"workload": "msi_diffusion_profiles: all drugs + indications of the MSI-shaped interactome, one batched fp64 "
                       "power iteration (BASELINE configs[2])",
           "nodes": n, "nnz": nnz, "seeds": ns, "sweeps": sweeps, "iterations_min_max": [int(it.min()), int(it.max())],
           "seconds_device": t_dev, "seeds_per_s": ns / t_dev,
           "effective_gedges_per_s": float(it.sum()) * nnz / t_dev / 1e9,
           "e2e": {"seconds": t_e2e, "seeds_per_s": ns / t_e2e, "d2h_bytes": int(x.numel() * 8),
                   "note": "solve + copy of all profiles to pinned host memory"},
           "roofline": {"kernel": "rw_step_kernel (one launch per sweep, all seeds)", "bound": "hbm",
                        "bytes_per_sweep_model": b_sweep, "achieved": gbs, "peak": peak, "unit": "GB/s",
                        "frac": gbs / peak, "peak_source": peak_src,
                        "note": "SURVEY 8(d) gather model nnz*12 + nnz*S*8 + 2*N*S*8 per sweep x sweeps / device time of "
                                "the whole solve (row scaling, dangling sums and convergence checks included)"},
           "check_vs_oracle": {"seeds": n_chk, "max_rel_diff": worst, "same_stopping_iteration": bool(it_same)},
           "cpu_baseline": {"kind": "port", "s_per_seed_1_core": t_cpu1, "seeds_per_s_1_core": 1.0 / t_cpu1,
                            "sample": "%d seeds, oracle/rwr_oracle.py (diffusion_profiles.py:29-90 restated)" % n_chk}}
    if with_cpu:
        import multiprocessing as mp
        workers = max(1, (os.cpu_count() or 2) // 2)                              # predict_drug.py:93
        per = 3
        jobs = [(seeds[w * per:(w + 1) * per],) for w in range(workers)]
        try:
            ctx = mp.get_context("spawn")                                         # never fork next to a CUDA context
            with ctx.Pool(workers) as pool:
                t0 = time.perf_counter()
                secs = pool.map(_rwr_cpu_worker, jobs)
                wall_pool = time.perf_counter() - t0
            out["cpu_baseline"].update({"workers": workers, "seeds_per_s_workers": workers * per / max(secs),
                                        "workers_sample": "%d workers x %d seeds each, slowest worker %.2f s (pool wall "
                                                          "%.1f s incl. start-up)" % (workers, per, max(secs), wall_pool)})
        except Exception as exc:  # noqa: BLE001
            out["cpu_baseline"]["workers_error"] = repr(exc)[:200]
    return out


# ------------------------------------------------------------------------------------------------------
# BASELINE configs[3]: R-MAT scale 20 (1M nodes / 32M edges), d 256: GCN training iteration + Gram top-k
# ------------------------------------------------------------------------------------------------------
def run_rmat20_gcn(args, scale=20, d=256, batch=2048, iters=5):
    import importlib
    import torch
    gdr = importlib.import_module("gcn-drug-repurposing_b200")
    from gdr_b200.graph import DeviceCSR
    dev = torch.device("cuda", 0)
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
    csr = build_rmat_csr(scale, 32, dev)
    n = csr.n_rows
    g = torch.Generator(device=dev).manual_seed(2)
    x = torch.randn(n, d, device=dev, generator=g)
    np.random.seed(1)
    torch.manual_seed(1)
    model = gdr.ResidualGraphConvolutionalNetwork(batch, n, num_layers=2, hidden_units=d, layer_decay=0.3).to(dev)
    opt = gdr.optim.Adam(model.parameters(), lr=3e-4, weight_decay=0)
    loss_fcn = gdr.GSS_loss(1.0).gss_loss
    from gdr_b200 import model as model_mod
    hoisted = bool(model_mod.HOIST_LAYER1)

    def iteration():
        idx = torch.randint(0, n, (batch,), device=dev, generator=g)
        h = model(x=x, adj=csr)
        loss = loss_fcn(embs=h, beta=0.1, index=idx)
        opt.zero_grad()
        loss.backward()
        opt.step()
        return loss

    def timed(k):
        for _ in range(3):
            iteration()
        torch.cuda.synchronize()
        DeviceCSR.timer = []
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            loss = iteration()
        e1.record()
        torch.cuda.synchronize()
        ev, DeviceCSR.timer = DeviceCSR.timer, None
        return e0.elapsed_time(e1) / k, ev, float(loss.item())

    model.hoist_layer1 = False
    ms_full, ev, loss = timed(iters)                      # all 6 SpMMs per iteration (the configuration of round 1)
    model.hoist_layer1 = hoisted
    ms_hoist, ev_h, _ = timed(iters) if hoisted else (ms_full, ev, loss)
    model.hoist_layer1 = False
    model.invalidate_layer1_cache()
    peak, peak_src = measured_peaks()
    plain = [(a.elapsed_time(b), nz, nr, dd) for a, b, nz, nr, dd, mul, add in ev if not mul and not add]
    ms_spmm = float(np.mean([t for t, *_ in plain])) if plain else None
    b_alg = spmm_alg_bytes(csr.nnz, n, d)
    out = {"workload": "rmat20_gcn: 2-layer GSS GCN training iteration (fwd + batch-%d GSS loss + bwd + Adam), R-MAT scale %d "
                       "edge-factor 32, d %d (BASELINE configs[3])" % (batch, scale, d),
           "nodes": n, "nnz": csr.nnz, "ms_per_iteration": ms_hoist, "ms_per_iteration_without_hoist": ms_full,
           "spmm_launches_per_iteration": len(ev) // iters, "loss": loss,
           "spmm": {"ms_per_launch_plain": ms_spmm, "alg_gb": b_alg / 1e9,
                    "achieved_gather_gbs": (b_alg / (ms_spmm * 1e-3) / 1e9) if ms_spmm else None, "peak_gbs": peak,
                    # no ncu capture at this size: the fraction is the compulsory one (every operand once)
                    "frac_compulsory": ((csr.nnz * 8 + 2 * n * d * 4 + (n + 1) * 4) / (ms_spmm * 1e-3) / 1e9 / peak)
                    if ms_spmm else None, "peak_source": peak_src,
                    "gedges_per_s": (csr.nnz / (ms_spmm * 1e-3) / 1e9) if ms_spmm else None,
                    "note": "SURVEY 8(d) gather model (35.7 GB at scale 20, d 256); X = %.2f GB > L2" % (n * d * 4 / 1e9)},
           "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9}
    # the other half of configs[3]: the kNN graph of gen_graph (Gram X X^T + row top-k, k = 5) at this size, on the
    # tensor cores (gdr_gram_topk_tc_f64: TF32 tiles + fused running top-C + exact fp64 re-rank, certified per row)
    helper = gdr.helper
    tf_peak = measured_tflops()
    tf32 = measure_tf32_tflops(dev)
    xd = x.double()
    helper.gram_topk(xd[:4096].contiguous(), 5, backend="tc")        # warm-up (module load, attribute calls)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    idx, val = helper.gram_topk(xd, 5, backend="tc")
    e1.record()
    torch.cuda.synchronize()
    secs = e0.elapsed_time(e1) * 1e-3
    fallback = int(helper.last_gram_topk_info["fallback_rows"].item())
    rows = torch.randint(0, n, (1024,), device=dev, generator=g)
    tv, ti = torch.topk(xd[rows] @ xd.T, 5, dim=1)                    # direct fp64 product on a row sample
    same = (torch.sort(ti, 1).values == torch.sort(idx[rows].long(), 1).values).all(1).double().mean().item()
    flops = 2.0 * n * n * d
    out["gram_topk"] = {"k": 5, "seconds": secs, "flops_model": flops, "tflops_tf32": flops / secs / 1e12,
                        "fallback_rows_fp64": fallback,
                        "peak_bf16_tflops_measured": tf_peak, "peak_tf32_tflops_measured": tf32,
                        "frac_of_tf32_peak": (flops / secs / 1e12) / tf32["sustained_tflops"],
                        "check_vs_fp64_product": {"sample_rows": 1024, "equal_index_sets": same,
                                                  "max_rel_value_diff": float(((tv - val[rows]).abs() / tv.abs()).max().item())},
                        "note": "helpers/helper.py:39,44 at N = 2^%d: TF32 tcgen05 Gram tiles (operand A resident in tensor "
                                "memory) with the running top-16 per row fused into the epilogue, exact fp64 re-rank of the "
                                "candidates, rows that cannot be certified recomputed in fp64; whole call timed "
                                "(prep + tensor pass + re-rank + fallback)" % scale}
    del xd, idx, val, tv, ti
    torch.cuda.empty_cache()
    # train.py --batch-size 0 / gss_loss(index=None): the N x N loss and its gradient in row panels, never stored
    with torch.no_grad():
        h = model(x=x, adj=csr)
    h = h.detach().requires_grad_(True)
    torch.cuda.synchronize()
    e0.record()
    loss_all = loss_fcn(embs=h, beta=0.1)
    loss_all.backward()
    e1.record()
    torch.cuda.synchronize()
    secs = e0.elapsed_time(e1) * 1e-3
    rows = torch.randint(0, n, (64,), device=dev, generator=g)
    hd = h.detach().double()
    gs = hd[rows] @ hd.T
    w = torch.where(gs > 0, -(gs - 0.1), torch.zeros_like(gs)) * (2.0 / (float(n) * float(n)))
    ref_rows = w @ hd
    err = (h.grad[rows].double() - ref_rows).abs()
    ok = ((err <= 1e-6) | (err <= 1e-5 * ref_rows.abs())).double().mean().item()
    out["full_batch_gss_loss"] = {"rows": n, "seconds_fwd_bwd": secs, "loss": float(loss_all.item()),
                                  "flops_model": 6.0 * n * n * d, "tflops_fp32_equivalent": 6.0 * n * n * d / secs / 1e12,
                                  "peak_tf32_tflops_measured": tf32["sustained_tflops"],
                                  "frac_of_tf32_peak_useful": 6.0 * n * n * d / secs / 1e12 / tf32["sustained_tflops"],
                                  "frac_of_tf32_peak_issued_3xtf32": 18.0 * n * n * d / secs / 1e12 / tf32["sustained_tflops"],
                                  "workspace_gb": gdr._ffi.lib().gdr_gss_loss_tc_workspace_bytes(n, d) / 1e9,
                                  "check_vs_fp64": {"gradient_rows": 64, "fraction_within_1e-5_rel_or_1e-6_abs": ok,
                                                    "max_abs_err": float(err.max().item()),
                                                    "max_abs_ref": float(ref_rows.abs().max().item())},
                                  "note": "modules/model.py:214-221 with index=None at N = 2^%d: Gram + loss/2dG epilogue and "
                                          "2dG.E per 1 GiB row panel (3xTF32 tcgen05), B x B never stored" % scale}
    del h, hd, gs, w, ref_rows
    del csr, x, model
    torch.cuda.empty_cache()
    return out


def measured_tflops():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f).get("bf16_tflops_sustained", 0.0)) or None
    return 1400.0


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
                "config": msi_config(n, int(cpu.nnz), MSI_D, iters_per_epoch),
                "cpu_baseline": {"value": value, "unit": "epochs/s", "cores": cores, "kind": "port", "sample": sample},
                "e2e": {"value": value, "unit": "epochs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
    # rmat_spmm: torch.sparse.mm on the host.  Preferred form: the SAME scale as the GPU arm — the matrix is generated on
    # this box's GPU by the generator the GPU arm uses (input preparation only; generating scale 24 on host cores takes
    # minutes), copied to the host, and every step multiplies one of 16 equal-nnz row blocks of it by the full X.
    try:
        line = _reference_rmat_same_scale(args, cores)
        if line is not None:
            return line
    except Exception as exc:                          # no usable GPU / not enough host memory: smaller graph below
        sys.stderr.write("reference arm: same-scale R-MAT unavailable (%s: %s), falling back to scale <= 22\n"
                         % (type(exc).__name__, exc))
    import importlib
    synth = importlib.import_module("gcn-drug-repurposing_b200.synth")
    from oracle import gcn_oracle
    import scipy.sparse as sp
    scale = min(args.rmat_scale, 22)                  # bounded sample: generating scale 24 on host cores takes minutes
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


def _reference_rmat_same_scale(args, cores, blocks=16):
    """torch.sparse.mm (COO, as modules/model.py:163 receives it) on the host cores over the GPU arm's own R-MAT matrix.
    Step i multiplies row block i mod `blocks` (blocks cut at equal nnz) by the full X; value = edges of the block / its
    time.  Returns None when no CUDA device is there to generate the graph."""
    import torch
    if not torch.cuda.is_available():
        return None
    import importlib
    synth = importlib.import_module("gcn-drug-repurposing_b200.synth")       # generator only; libgdr.so is never loaded
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    n, d = 1 << args.rmat_scale, args.rmat_d
    # the GPU arm's matrix (build_rmat_csr: same generator, seeds and (A+I) normalisation), built with torch ops only
    r, c = synth.rmat_coo(args.rmat_scale, args.rmat_edge_factor, 1, dev)
    keys = torch.unique(r * n + c)
    del r, c
    vals = 1.0 - torch.rand(keys.numel(), generator=torch.Generator(device=dev).manual_seed(2), device=dev,
                            dtype=torch.float64)
    keys, inv = torch.unique(torch.cat([keys, torch.arange(n, device=dev) * (n + 1)]), return_inverse=True)
    vals = torch.zeros(keys.numel(), dtype=torch.float64, device=dev).index_add_(
        0, inv, torch.cat([vals, torch.ones(n, dtype=torch.float64, device=dev)]))
    del inv
    rows_d = torch.div(keys, n, rounding_mode="floor")
    col_d = keys - rows_d * n
    del keys
    dinv = torch.zeros(n, dtype=torch.float64, device=dev).index_add_(0, rows_d, vals).pow_(-0.5)
    val = (dinv[rows_d] * vals * dinv[col_d]).float().cpu()
    rowptr = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    rowptr[1:] = torch.cumsum(torch.bincount(rows_d, minlength=n), 0)
    rowptr, col = rowptr.cpu(), col_d.cpu()
    nnz = int(val.numel())
    del rows_d, col_d, vals, dinv
    torch.cuda.empty_cache()
    x = torch.randn(n, d, generator=torch.Generator().manual_seed(2))
    cuts = torch.searchsorted(rowptr, torch.arange(blocks + 1, dtype=torch.int64) * nnz // blocks).tolist()
    cuts[0], cuts[-1] = 0, n

    def block(i):
        r0, r1 = cuts[i % blocks], cuts[i % blocks + 1]
        p0, p1 = int(rowptr[r0]), int(rowptr[r1])
        rows = torch.repeat_interleave(torch.arange(r1 - r0, dtype=torch.int64), rowptr[r0 + 1:r1 + 1] - rowptr[r0:r1])
        coo = torch.sparse_coo_tensor(torch.stack([rows, col[p0:p1]]), val[p0:p1], (r1 - r0, n), is_coalesced=True)
        return coo, p1 - p0

    for i in range(min(args.warmup, 2)):
        a, _ = block(blocks - 1 - i)
        torch.sparse.mm(a, x)
    times, edges = [], []
    for i in range(args.steps):
        a, e = block(i)
        t0 = time.perf_counter()
        torch.sparse.mm(a, x)
        times.append(time.perf_counter() - t0)
        edges.append(e)
    value = float(np.sum(edges)) / float(np.sum(times)) / 1e9
    sample = ("step i = row block i mod %d (blocks cut at equal nnz, %.1f M edges each) of the scale-%d matrix times the "
              "full X, torch.sparse.mm on %d host threads; graph generated on the GPU by the GPU arm's generator and "
              "copied to the host (input preparation, untimed)" % (blocks, nnz / blocks / 1e6, args.rmat_scale, cores))
    return {"impl": "reference", "metric": "spmm_gedges_per_s", "value": value, "unit": "GEdges/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": float(np.mean(times)) * 1e3, "full_product_ms_extrapolated": nnz / value / 1e6,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "rmat_spmm: Y = A_hat X, R-MAT scale %d edge-factor %d d %d, rows partitioned over the "
                                   "GPUs, exchange of the referenced X rows / partial Y rows + local SpMM per step "
                                   "(BASELINE configs[4])" % (args.rmat_scale, args.rmat_edge_factor, d),
                       "nodes": n, "nnz": nnz, "d": d, "parallelism": "host threads (torch.sparse.mm)"},
            "cpu_baseline": {"value": value, "unit": "GEdges/s", "cores": cores, "kind": "port", "sample": sample},
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
        if not args.no_rwr:
            line["rwr"] = run_rwr(args, with_cpu=not args.no_cpu_baseline)
        if not args.no_rmat20:
            line["rmat20_gcn"] = run_rmat20_gcn(args)
        if not args.no_cpu_baseline:
            import torch
            cpu = CpuReferenceGCN(a, feats)
            cpu.set_beta(float(beta))           # same beta as the GPU arm; the percentile is not part of the step
            flat = [b for ep in host_batches for b in ep]
            cpu.iteration(flat[0])              # warm-up
            times = []
            for b in flat[1:1 + args.cpu_iters]:
                t0 = time.perf_counter()
                cpu.iteration(b)
                times.append(time.perf_counter() - t0)
            ipe = line["config"]["iters_per_epoch"]
            med = float(np.median(times))
            line["cpu_baseline"] = {"value": 1.0 / (med * ipe), "unit": "epochs/s",
                                    "cores": torch.get_num_threads(), "kind": "port",
                                    "sample": "median of %d of the %d iterations of one epoch (after 1 warm-up), %.1f s of CPU "
                                              "work, torch CPU restatement of modules/model.py + train.py "
                                              "(oracle/gcn_oracle.py)" % (len(times), ipe, float(np.sum(times)))}
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
               "roofline": {"kernel": "spmm_warp_row_kernel<float>", "bound": "hbm",
                            "achieved": (line["frac"] if line["frac"] is not None else line["frac_compulsory"]) * peak,
                            "peak": peak, "unit": "GB/s",
                            "frac": line["frac"] if line["frac"] is not None else line["frac_compulsory"],
                            "achieved_basis": "ncu DRAM bytes per launch / live launch time" if line["frac"] is not None
                            else "compulsory bytes / live launch time",
                            "achieved_gather_gbs": line["achieved_gather_gbs"],
                            "traffic": line["traffic"], "peak_source": src},
               "spmm_rmat": line}
        print(json.dumps(out), flush=True)
        return 0
    line = run_rmat_spmm_dist(args)
    if line is not None:
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
