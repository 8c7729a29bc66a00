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
