"""torchrun worker of tests/test_gpu_dist.py: one rank per GPU (2 ranks), NCCL + symmetric memory.
Checks every transport / option of the partitioned SpMM against the 1-GPU product, with X changed between
consecutive steps (the exchange buffers must not leak rows of the previous step), and the two NVSwitch multicast
primitives.  Prints `DIST_GPU_OK <rank>` on success; any assertion failure exits non-zero."""
import datetime
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import gdr_b200 as gdr  # noqa: E402
from gdr_b200 import dist as gdist  # noqa: E402


def close(y, ref, what):
    err = (y - ref).abs()
    ok = (err <= 1e-6) | (err <= 1e-5 * ref.abs())
    frac = float(ok.double().mean().item())
    # different fp32 summation order than the 1-GPU kernel on hub rows: strict on >= 99.9 %, all within 4e-7*max|y|
    assert frac >= 0.999, f"{what}: only {frac:.6f} of the entries within 1e-5 rel / 1e-6 abs"
    assert bool((ok | (err <= 4e-7 * ref.abs().max())).all().item()), f"{what}: max abs err {float(err.max()):.3e}"


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    scale, d = 16, 128
    full = bench.build_rmat_csr(scale, 16, dev)
    n = full.n_rows
    deg = (full.rowptr[1:] - full.rowptr[:-1]).to(torch.int64)
    row_of = torch.repeat_interleave(torch.arange(n, device=dev), deg)
    hist, bin_size = gdist.coarse_histogram(row_of, full.col.to(torch.int64), n, min(1024, n))
    bounds = gdist.lshape_bounds(hist.cpu().numpy(), bin_size, n, world, row_weight=6.0)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    g = torch.Generator(device=dev).manual_seed(2)
    xs = [torch.randn(n, d, device=dev, generator=g) for _ in range(3)]
    refs = [full.spmm(x)[lo:hi].clone() for x in xs]
    p2p = gdist.p2p_available(dev)
    variants = [("nccl", dict(transport="nccl"))]
    if p2p:
        variants += [("p2p", dict(transport="p2p")),
                     ("p2p+hub_first", dict(transport="p2p", hub_first=True)),
                     ("p2p+split_pull", dict(transport="p2p", split_pull="force")),
                     ("p2p+hub_first+split_pull", dict(transport="p2p", hub_first=True, split_pull="force")),
                     ("p2p+x_priority", dict(transport="p2p", x_priority=1)),
                     ("p2p+fused_push", dict(transport="p2p", fused_push=True)),
                     ("p2p+fused_push+split_pull", dict(transport="p2p", fused_push=True, split_pull="force"))]
    for name, kw in variants:
        part = gdist.LShapedSpMM.from_full(full, bounds, rank, world, d, **kw)
        if "split_pull" in kw and rank > 0:
            assert part.split_pull, "split_pull was requested but not set up"
            assert part.pull_early.nnz + (part.pull_late.nnz if part.pull_late is not None else 0) + \
                (part.final.nnz - part.n_precv if part.final is not None else 0) == part.pull.nnz
        for step, (x, ref) in enumerate(zip(xs, refs)):          # X changes between consecutive steps
            y = part.spmm(x[lo:hi])
            close(y, ref, f"{name} step {step} rank {rank}")
            first = y.clone()
        assert torch.equal(part.spmm(xs[-1][lo:hi]), first), f"{name}: two steps on the same input differ"
        # the row-sliced push product (hub_first) is the same kernel on a row range: identical bits
        if "fused_push" in kw:
            assert part.fused_push and part.psend.numel() == 0, "fused_push was requested but not set up"
        if part.n_psend > 1 and not part.fused_push:
            part.push_product()
            whole = part.psend.clone()
            mid = part.n_psend // 3
            part.psend.zero_()
            part._push_product_rows(0, mid)
            part._push_product_rows(mid, part.n_psend)
            assert torch.equal(part.psend, whole), f"{name}: sliced push product differs"
        torch.cuda.synchronize()
        dist.barrier()
        del part
    # ---- cyclic virtual-block layout (hub blocks spread over all ranks; renumbered vertex order) -----------------
    vb = gdist.lshape_bounds(hist.cpu().numpy(), bin_size, n, world * 3, row_weight=6.0)
    for name, kw in variants[:2]:
        part = gdist.LShapedSpMM.from_full_cyclic(full, vb, rank, world, d, **kw)
        sel = part.rows_old
        assert sum(part.xrecv_splits) > 0 and part.xrecv_splits[rank] == 0
        for step, x in enumerate(xs):
            close(part.spmm(x[sel]), full.spmm(x)[sel], f"cyclic {name} step {step} rank {rank}")
        torch.cuda.synchronize()
        dist.barrier()
        del part
    # ---- the transposed product (backward aggregation A_hat^T G): a second partition of the explicit transpose, with
    # bounds skewed by rank weights (what the measured rebalancing of bench.py produces) -------------------------
    full_t = full.transpose()
    deg_t = (full_t.rowptr[1:] - full_t.rowptr[:-1]).to(torch.int64)
    row_t = torch.repeat_interleave(torch.arange(n, device=dev), deg_t)
    hist_t, bs_t = gdist.coarse_histogram(row_t, full_t.col.to(torch.int64), n, min(1024, n))
    w = gdist.rebalance_weights(np.ones(world), np.linspace(1.0, 2.0, world))
    bounds_t = gdist.lshape_bounds(hist_t.cpu().numpy(), bs_t, n, world, row_weight=6.0, rank_weights=w)
    lt, ht = int(bounds_t[rank]), int(bounds_t[rank + 1])
    for name, kw in variants[:1] + variants[-2:-1]:
        part = gdist.LShapedSpMM.from_full(full_t, bounds_t, rank, world, d, **kw)
        for step, x in enumerate(xs[:2]):
            close(part.spmm(x[lt:ht]), full_t.spmm(x)[lt:ht], f"transposed {name} step {step} rank {rank}")
        torch.cuda.synchronize()
        dist.barrier()
        del part
    rp = gdist.RowPartitionedSpMM.from_full(full, gdist.nnz_balanced_bounds(full.rowptr.cpu().numpy(), world), rank,
                                            world, d)
    b2 = rp.bounds
    for x in xs[:2]:
        close(rp.spmm(x[b2[rank]:b2[rank + 1]]), full.spmm(x)[b2[rank]:b2[rank + 1]], f"row partition rank {rank}")
    # ---- NVSwitch multicast primitives (gdr_mc_broadcast_f32 / gdr_mc_reduce_add_f32) ------------------------
    if p2p:
        import torch.distributed._symmetric_memory as symm
        cnt = 1 << 16
        t = symm.empty(2 * cnt, dtype=torch.float32, device=dev)
        h = symm.rendezvous(t, dist.group.WORLD.group_name)
        mc = int(getattr(h, "multicast_ptr", 0) or 0)
        if mc:
            L = gdr._ffi.lib()
            t.zero_()
            torch.cuda.synchronize()
            h.barrier(channel=0)
            src = torch.arange(cnt, device=dev, dtype=torch.float32) + 1000.0 * rank
            if rank == 0:                                          # rank 0's block reaches every GPU
                gdr._ffi.check(L.gdr_mc_broadcast_f32(gdr._ffi.ptr(src), mc, cnt, 0, gdr._ffi.stream()), "mc_broadcast")
            torch.cuda.synchronize()
            h.barrier(channel=1)
            assert torch.equal(t[:cnt], torch.arange(cnt, device=dev, dtype=torch.float32)), "mc_broadcast"
            t[cnt:].copy_(src)                                     # every rank's partial buffer
            torch.cuda.synchronize()
            h.barrier(channel=2)
            y = torch.ones(cnt, device=dev)
            gdr._ffi.check(L.gdr_mc_reduce_add_f32(gdr._ffi.ptr(y), mc + cnt * 4, cnt, 0, gdr._ffi.stream()),
                           "mc_reduce_add")
            torch.cuda.synchronize()
            exp = 1.0 + world * torch.arange(cnt, device=dev, dtype=torch.float32) + 1000.0 * sum(range(world))
            assert torch.equal(y, exp), "mc_reduce_add"
            h.barrier(channel=3)
            print(f"multicast primitives ok on rank {rank}", flush=True)
        else:
            print("no multicast mapping on this box: multimem primitives not exercised", flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    print(f"DIST_GPU_OK {rank} variants={[v for v, _ in variants]}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
