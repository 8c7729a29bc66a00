"""Timing of gen_graph's kNN selection (helpers/helper.py:39,44) on one B200: tensor-core path
(gdr_gram_topk_tc_f64) vs the exact fp64 SIMT kernel (gdr_gram_topk_f64), CUDA events, after warm-up.

    python profiles/bench_gram_topk.py [msi|rmat20|both] [--f64]   ->  one JSON line per case

msi:    N = 29,960, d = 128, k = 5 (SURVEY 8 "MSI-kNN")
rmat20: N = 1,048,576, d = 256, k = 5 (BASELINE configs[3] "Gram top-k"); the fp64 kernel is timed on a row sample only.
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gdr_b200 as gdr  # noqa: E402

PEAKS = {}
try:
    PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    pass


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def case(name, n, d, k, reps, f64_rows):
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.randn(n, d, device="cuda", dtype=torch.float64, generator=g)
    x *= torch.empty(n, 1, device="cuda", dtype=torch.float64).uniform_(0.5, 3.0, generator=g)
    ms = timed(lambda: gdr.helper.gram_topk(x, k, backend="tc"), reps)
    idx, val = gdr.helper.gram_topk(x, k, backend="tc")
    nfb = int(gdr.helper.last_gram_topk_info["fallback_rows"].item())
    flops = 2.0 * n * n * d
    out = {"case": name, "n": n, "d": d, "k": k, "ms_tc": ms, "tflops_tf32": flops / ms / 1e9,
           "fallback_rows": nfb, "alg_flops": flops}
    if PEAKS.get("bf16_tflops"):
        # TF32 dense peak of the tcgen05 pipe is half the bf16 one (same datapath, 2x the operand bits)
        out["peak_tf32_tflops_assumed"] = PEAKS["bf16_tflops"] / 2
        out["frac_of_tf32_peak"] = out["tflops_tf32"] / (PEAKS["bf16_tflops"] / 2)
    # exact kernel on a row sample (all rows where affordable); same columns, so time scales with the rows
    m = min(n, f64_rows)
    xs = x if m == n else None
    if xs is not None:
        ms64 = timed(lambda: gdr.helper.gram_topk(x, k, backend="f64"), 1)
        i64, v64 = gdr.helper.gram_topk(x, k, backend="f64")
        out["ms_f64"] = ms64
        out["equal_values"] = bool(torch.equal(v64, val))
        out["equal_index_rows"] = float((i64 == idx).all(dim=1).double().mean().item())
    else:
        # check a sample of rows against a direct fp64 product
        rows = torch.randint(0, n, (m,), device="cuda", generator=g)
        t0 = time.time()
        s = x[rows] @ x.T
        tv, ti = torch.topk(s, k, dim=1)
        torch.cuda.synchronize()
        out["sample_rows"] = m
        out["sample_check_s"] = time.time() - t0
        out["sample_equal_sets"] = float((torch.sort(ti, 1).values == torch.sort(idx[rows].long(), 1).values).all(1).double().mean().item())
        out["sample_max_rel_val_diff"] = float(((tv - val[rows]).abs() / tv.abs()).max().item())
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "both"
    if which in ("msi", "both"):
        case("msi_knn", 29960, 128, 5, 5, 29960)
    if which in ("rmat20", "both"):
        case("rmat20_d256", 1 << 20, 256, 5, 1, 4096)
