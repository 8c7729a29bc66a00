"""CPU tests (gloo, world_size 2 and 3) of the host-side logic of the row-partitioned SpMM (dist.py): nnz-balanced
bounds, the compressed-exchange plan (which rows every rank needs from every owner, column renumbering), the
all-to-all of index lists and of the packed rows, and the own/remote column split.  The device product itself is
replaced by scipy here (the CUDA kernel is covered by the -m gpu tests and checked against the 1-GPU product inside
bench.py at N > 1)."""
import os
import socket

import numpy as np
import pytest
import scipy.sparse as sp
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _matrix(n, seed):
    rng = np.random.default_rng(seed)
    m = sp.random(n, n, density=0.01, random_state=rng, format="lil", dtype=np.float64)
    m[0, rng.choice(n, size=n // 2, replace=False)] = 1.0      # a hub row: nnz balance != row balance
    m = (m + sp.identity(n)).tocsr().astype(np.float32)
    m.sort_indices()
    return m


def _worker(rank, world, port, n, d, overlap, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import gdr_b200  # noqa: F401
    from gdr_b200 import dist as gdist
    from gdr_b200.graph import DeviceCSR
    m = _matrix(n, 3)
    x = np.random.default_rng(4).standard_normal((n, d)).astype(np.float32)
    bounds = gdist.nnz_balanced_bounds(m.indptr, world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    blk = m[lo:hi]
    local = DeviceCSR(torch.from_numpy(blk.indptr.astype(np.int32)), torch.from_numpy(blk.indices.astype(np.int32)),
                      torch.from_numpy(blk.data), hi - lo, n)
    part = gdist.RowPartitionedSpMM(local, bounds, rank, world, d, overlap=overlap)
    # the exchange step with the device kernels replaced by torch/scipy equivalents
    part.x_own.copy_(torch.from_numpy(x[lo:hi]))
    part.sendbuf.copy_(part.x_own[part.send_idx])                      # = gdr_gather_rows_f32
    dist.all_to_all_single(part.x_recv, part.sendbuf, part.recv_splits, part.send_splits)
    xb = part.xbuf.numpy()

    def to_sp(c):
        return sp.csr_matrix((c.val.numpy(), c.col.numpy(), c.rowptr.numpy()), shape=(c.n_rows, c.n_cols))

    if overlap:
        y = to_sp(part.csr_own) @ xb + to_sp(part.csr_rem) @ xb
        assert part.csr_own.nnz + part.csr_rem.nnz == blk.nnz
        assert int(part.csr_own.col.max()) < part.n_own                # own split never touches received rows
    else:
        y = to_sp(part.csr_all) @ xb
    exp = blk @ x
    ok = np.allclose(y, exp, rtol=1e-5, atol=1e-5)
    fewer = part.recv_rows <= n - (hi - lo)
    out.put((rank, bool(ok), bool(fewer), part.recv_rows, [int(b) for b in bounds]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,overlap", [(2, True), (2, False), (3, True)])
def test_row_partitioned_exchange_gloo(world, overlap):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 600, 8, overlap, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _, _, _ in res), res
    assert all(fewer for _, _, fewer, _, _ in res)
    bounds = res[0][4]
    assert bounds[0] == 0 and bounds[-1] == 600 and all(b1 >= b0 for b0, b1 in zip(bounds, bounds[1:]))


def test_nnz_balanced_bounds_balance_nnz_not_rows():
    import gdr_b200  # noqa: F401
    from gdr_b200 import dist as gdist
    m = _matrix(2000, 1)
    b = gdist.nnz_balanced_bounds(m.indptr, 4)
    per = np.diff(m.indptr[b])
    assert per.sum() == m.nnz and per.max() < 1.6 * m.nnz / 4
    assert gdist.nnz_balanced_bounds(m.indptr, 1).tolist() == [0, 2000]


def test_plan_exchange_renumbering():
    import gdr_b200  # noqa: F401
    from gdr_b200 import dist as gdist
    col = torch.tensor([0, 5, 9, 5, 3, 7, 8, 1], dtype=torch.int64)
    need, new_col, own = gdist.plan_exchange(col, [0, 4, 8, 10], rank=1)     # own block = columns 4..7
    assert [t.tolist() for t in need] == [[0, 1, 3], [], [0, 1]]              # owner-local ids
    assert own.tolist() == [False, True, False, True, False, True, False, False]
    # compact buffer: [own 4 rows | from rank 0: cols 0,1,3 | from rank 2: cols 8,9]
    assert new_col.tolist() == [4, 1, 8, 1, 6, 3, 7, 5]


# ---- L-shaped partition (push blocks above the diagonal) -------------------------------------------------
def _skewed_matrix(n, seed):
    """Rows and columns both skewed towards low indices (R-MAT-like), plus the identity."""
    rng = np.random.default_rng(seed)
    m = 40 * n
    r = np.minimum((rng.random(m) ** 3 * n).astype(np.int64), n - 1)
    c = np.minimum((rng.random(m) ** 3 * n).astype(np.int64), n - 1)
    a = sp.coo_matrix((rng.random(m).astype(np.float32) + 0.1, (r, c)), shape=(n, n)).tocsr()
    a = (a + sp.identity(n, dtype=np.float32)).tocsr().astype(np.float32)
    a.sort_indices()
    return a


def _lshape_worker(rank, world, port, n, d, out, cyclic_m=0):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import gdr_b200  # noqa: F401
    from gdr_b200 import dist as gdist
    from gdr_b200.graph import DeviceCSR
    m = _skewed_matrix(n, 5)
    x = np.random.default_rng(6).standard_normal((n, d)).astype(np.float32)
    coo = m.tocoo()
    hist, size = gdist.coarse_histogram(torch.from_numpy(coo.row.astype(np.int64)),
                                        torch.from_numpy(coo.col.astype(np.int64)), n, 64)
    if cyclic_m:
        # cyclic ownership of world * cyclic_m virtual blocks: the object works in a renumbered vertex order
        vb = gdist.lshape_bounds(hist.numpy(), size, n, world * cyclic_m, row_weight=2.0)
        full = DeviceCSR(torch.from_numpy(m.indptr.astype(np.int32)), torch.from_numpy(m.indices.astype(np.int32)),
                         torch.from_numpy(m.data), n, n)
        part = gdist.LShapedSpMM.from_full_cyclic(full, vb, rank, world, d)
        rows_old = part.rows_old.numpy()
        bounds = part.bounds
        blk = m[rows_old]
        x_mine = x[rows_old]
    else:
        bounds = gdist.lshape_bounds(hist.numpy(), size, n, world, row_weight=2.0)
        lo, hi = int(bounds[rank]), int(bounds[rank + 1])
        blk = m[lo:hi]
        x_mine = x[lo:hi]
        local = DeviceCSR(torch.from_numpy(blk.indptr.astype(np.int32)), torch.from_numpy(blk.indices.astype(np.int32)),
                          torch.from_numpy(blk.data), hi - lo, n)
        part = gdist.LShapedSpMM(local, bounds, rank, world, d)

    def to_sp(c):
        return sp.csr_matrix((c.val.numpy(), c.col.numpy(), c.rowptr.numpy()), shape=(c.n_rows, c.n_cols))

    # one step with the device kernels replaced by torch/scipy equivalents, same order as LShapedSpMM.spmm
    part.x_own.copy_(torch.from_numpy(x_mine))
    part.xsend.copy_(part.x_own[part.send_idx])                                    # = gdr_gather_rows_f32
    dist.all_to_all_single(part.x_recv, part.xsend, part.xrecv_splits, part.xsend_splits)
    if part.n_psend:
        part.psend.copy_(torch.from_numpy(to_sp(part.push) @ part.x_own.numpy()))
    part.exchange_partials()
    y = to_sp(part.pull) @ part.xbuf.numpy()
    if part.n_precv:
        y = y + to_sp(part.add)[:, :part.n_precv] @ part.precv.numpy()
    exp = blk @ x
    ok = np.allclose(y, exp, rtol=1e-5, atol=1e-5)
    # bookkeeping: every entry of the matrix is multiplied exactly once, somewhere
    nnz_all = torch.tensor([part.pull.nnz + part.push.nnz], dtype=torch.int64)
    dist.all_reduce(nnz_all)
    # the pull part never references columns of higher ranks; the push part only own columns
    own_cols_only = part.push.nnz == 0 or int(part.push.col.max()) < part.n_own
    no_recv_from_higher = all(c == 0 for c in part.xrecv_splits[rank:]) or bool(cyclic_m)
    no_partials_from_lower = all(c == 0 for c in part.precv_splits[:rank + 1]) or bool(cyclic_m)
    if cyclic_m:   # both directions between every pair of ranks; nothing to or from itself
        own_cols_only = own_cols_only and part.xrecv_splits[rank] == 0 and part.precv_splits[rank] == 0
    out.put((rank, bool(ok), int(nnz_all.item()) == m.nnz, bool(own_cols_only and no_recv_from_higher
                                                                  and no_partials_from_lower),
             part.recv_rows, part.pull.nnz + part.push.nnz, [int(b) for b in bounds]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_lshaped_partition_gloo(world):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    n = 700
    procs = [ctx.Process(target=_lshape_worker, args=(r, world, port, n, 8, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(r[1] for r in res), res                    # product equals the row block of the full product
    assert all(r[2] for r in res), res                    # no entry lost or duplicated
    assert all(r[3] for r in res), res                    # direction of both exchanges
    bounds = res[0][6]
    assert bounds[0] == 0 and bounds[-1] == n and all(b1 >= b0 for b0, b1 in zip(bounds, bounds[1:]))
    loads = [r[5] for r in sorted(res)]
    assert max(loads) < 1.5 * sum(loads) / world, loads  # L-shapes, not row blocks, carry equal nnz


@pytest.mark.parametrize("world,m", [(2, 3), (3, 2)])
def test_lshaped_partition_cyclic_virtual_blocks_gloo(world, m):
    """The cyclic virtual-block layout (hub blocks spread over all ranks): same product, every entry multiplied once."""
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    n = 700
    procs = [ctx.Process(target=_lshape_worker, args=(r, world, port, n, 8, out, m)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(r[1] for r in res), res
    assert all(r[2] for r in res), res
    assert all(r[3] for r in res), res
    bounds = res[0][6]
    assert bounds[0] == 0 and bounds[-1] == n
    recv = [r[4] for r in sorted(res)]
    assert min(recv) > 0                                   # every rank receives rows (no rank is a pure sender)


def test_cyclic_vblock_layout_is_a_permutation():
    import gdr_b200  # noqa: F401
    from gdr_b200 import dist as gdist
    vb = np.array([0, 3, 3, 10, 14, 20, 21], dtype=np.int64)        # 6 virtual blocks (one empty), 3 ranks
    order, new_vb, rb = gdist.cyclic_vblock_layout(vb, 3)
    assert order.tolist() == [0, 3, 1, 4, 2, 5]
    assert new_vb.tolist() == [0, 3, 7, 7, 13, 20, 21] and rb.tolist() == [0, 7, 13, 21]
    perm, inv = gdist.cyclic_permutation(vb, order, torch.device("cpu"))
    assert sorted(perm.tolist()) == list(range(21)) and torch.equal(perm[inv], torch.arange(21))
    assert perm[:7].tolist() == [0, 1, 2, 10, 11, 12, 13]
    with pytest.raises(ValueError):
        gdist.cyclic_vblock_layout(vb[:-1], 3)


def test_lshape_bounds_balances_l_shapes():
    import gdr_b200  # noqa: F401
    from gdr_b200 import dist as gdist
    m = _skewed_matrix(4000, 2).tocoo()
    hist, size = gdist.coarse_histogram(torch.from_numpy(m.row.astype(np.int64)),
                                        torch.from_numpy(m.col.astype(np.int64)), 4000, 128)
    assert int(hist.sum()) == m.nnz
    b = gdist.lshape_bounds(hist.numpy(), size, 4000, 4, row_weight=0.0)
    assert b[0] == 0 and b[-1] == 4000 and (np.diff(b) > 0).all()
    blk_r = np.searchsorted(b, m.row, side="right") - 1
    blk_c = np.searchsorted(b, m.col, side="right") - 1
    load = np.bincount(np.maximum(blk_r, blk_c), minlength=4)      # block (r, s) is computed by max(r, s)
    assert load.sum() == m.nnz and load.max() < 1.25 * m.nnz / 4, load
    assert gdist.lshape_bounds(hist.numpy(), size, 4000, 1).tolist() == [0, 4000]


def test_step_model_and_refinement():
    """The timeline model's volumes track the real ones, and the descent never worsens its own objective."""
    import gdr_b200  # noqa: F401
    from gdr_b200 import dist as gdist
    n, p = 6000, 4
    m = _skewed_matrix(n, 7).tocoo()
    hist, size = gdist.coarse_histogram(torch.from_numpy(m.row.astype(np.int64)),
                                        torch.from_numpy(m.col.astype(np.int64)), n, 200)
    b0 = gdist.lshape_bounds(hist.numpy(), size, n, p, row_weight=2.0)
    t0, det = gdist.lshape_step_model(hist.numpy(), size, n, [int(b // size) for b in b0], 128,
                                       consts=dict(volume_scale=1.0))    # smooth test matrix: Poisson is exact
    blk_r = np.searchsorted(b0, m.row, side="right") - 1
    blk_c = np.searchsorted(b0, m.col, side="right") - 1
    sent_x = np.zeros(p)
    for r in range(p):
        for s in range(r):
            sent_x[s] += len(np.unique(m.col[(blk_r == r) & (blk_c == s)]))
    assert np.allclose(det["sent_x"], sent_x, rtol=0.15, atol=30), (det["sent_x"], sent_x)
    assert abs(det["nnz_pull"].sum() + det["nnz_push"].sum() - m.nnz) < 1e-6 * m.nnz
    b1, t1 = gdist.refine_lshape_bounds(hist.numpy(), size, n, b0, 128, consts=dict(volume_scale=1.0))
    assert t1 <= t0 * (1 + 1e-9) and b1[0] == 0 and b1[-1] == n and (np.diff(b1) > 0).all()


def test_lshape_bounds_follow_rank_weights():
    """rank_weights (the measured-finish-time feedback of bench.py) shrink the region of a heavy rank and keep the
    bounds a valid cover; rebalance_weights is normalised and monotone in the finish times."""
    from gdr_b200 import dist as gdist
    rng = np.random.default_rng(0)
    nb, n = 64, 64 * 100
    hist = rng.integers(50, 200, size=(nb, nb)).astype(np.float64)
    hist[:8, :] *= 20
    hist[:, :8] *= 20                                            # hub rows / columns, as on R-MAT
    base = gdist.lshape_bounds(hist, 100, n, 4)
    w = gdist.rebalance_weights(np.ones(4), [1.0, 1.0, 1.0, 2.0])
    assert abs(w.mean() - 1.0) < 1e-12 and w[3] > w[0] and np.all(np.diff(w[:3]) == 0)
    skew = gdist.lshape_bounds(hist, 100, n, 4, rank_weights=w)
    for b in (base, skew):
        assert b[0] == 0 and b[-1] == n and np.all(np.diff(b) >= 0)
    assert (skew[4] - skew[3]) < (base[4] - base[3])             # the slow last rank gets fewer rows
    same = gdist.lshape_bounds(hist, 100, n, 4, rank_weights=np.ones(4))
    assert np.array_equal(same, base)
