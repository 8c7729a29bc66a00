"""GPU tests of the multi-GPU data path (csrc/p2p.cu + dist.py): the peer-store kernel against torch.index_select
on ONE GPU (local destination pointers: covers CTA dealing, idx == NULL ranges, ragged and empty destinations), and
a 2-rank torchrun run of every transport / option of the partitioned SpMM against the 1-GPU product (skipped when
the box has fewer than 2 GPUs)."""
import ctypes
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("d,counts,n_ctas", [(128, [1000, 0, 37, 5000], 0), (256, [1, 1, 1], 3), (4, [70000], 7),
                                            (512, [300, 4096, 2], 132), (128, [0, 0, 9], 1)])
def test_push_rows_local_destinations_match_index_select(gdr, d, counts, n_ctas):
    dev = torch.device("cuda", 0)
    L = gdr._ffi.lib()
    g = torch.Generator(device=dev).manual_seed(d + len(counts))
    n_src = 9000
    src = torch.randn(n_src, d, device=dev, generator=g)
    k = len(counts)
    idx = [torch.randint(0, n_src, (c,), device=dev, generator=g) if (i % 2 == 0) else None
           for i, c in enumerate(counts)]                               # odd destinations: contiguous ranges (idx NULL)
    offs = [int(torch.randint(0, n_src - c, (1,)).item()) if (c and idx[i] is None) else 0 for i, c in enumerate(counts)]
    dst = [torch.full((max(c, 1), d), float("nan"), device=dev) for c in counts]
    arr, i64 = ctypes.c_void_p * k, ctypes.c_int64 * k
    rc = L.gdr_push_rows_f32(src.data_ptr(), d, k, arr(*[t.data_ptr() for t in dst]),
                             arr(*[(t.data_ptr() if t is not None and t.numel() else None) for t in idx]),
                             i64(*offs), i64(*counts), n_ctas, gdr._ffi.stream())
    gdr._ffi.check(rc, "push_rows")
    torch.cuda.synchronize()
    for i, c in enumerate(counts):
        if c == 0:
            assert torch.isnan(dst[i]).all()                            # untouched
            continue
        exp = src.index_select(0, idx[i]) if idx[i] is not None else src[offs[i]:offs[i] + c]
        assert torch.equal(dst[i][:c], exp), f"destination {i}"


def test_push_rows_argument_errors(gdr):
    L = gdr._ffi.lib()
    arr, i64 = ctypes.c_void_p * 1, ctypes.c_int64 * 1
    x = torch.zeros(4, 6, device="cuda")
    assert L.gdr_push_rows_f32(x.data_ptr(), 6, 1, arr(x.data_ptr()), arr(None), i64(0), i64(1), 0, None) == -2   # d % 4
    assert L.gdr_push_rows_f32(x.data_ptr(), 8, 17, arr(x.data_ptr()), arr(None), i64(0), i64(1), 0, None) == -2  # > 16
    assert L.gdr_push_rows_f32(x.data_ptr(), 8, 1, arr(None), arr(None), i64(0), i64(1), 0, None) == -1           # null dst


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (one rank per GPU)")
def test_partitioned_spmm_two_ranks_against_one_gpu_product():
    """LShapedSpMM (nccl / p2p / hub_first / split_pull), RowPartitionedSpMM and the multimem primitives."""
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29617", os.path.join(ROOT, "tests", "dist_gpu_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    tail = (out.stdout[-3000:] + "\n" + out.stderr[-3000:])
    assert out.returncode == 0, tail
    assert "DIST_GPU_OK 0" in out.stdout and "DIST_GPU_OK 1" in out.stdout, tail
