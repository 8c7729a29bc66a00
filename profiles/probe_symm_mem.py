"""Probe: does torch's symmetric memory (peer pointers + NVLS multicast pointer) come up on this box?
    torchrun --nproc-per-node N profiles/probe_symm_mem.py"""
import os
import torch
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1 << 20, dtype=torch.float32, device=dev)
    h = symm.rendezvous(t, dist.group.WORLD.group_name)
    info = dict(rank=rank, buffer_ptrs=[hex(p) for p in h.buffer_ptrs], multicast_ptr=hex(getattr(h, "multicast_ptr", 0) or 0),
                signal_pads=len(h.signal_pad_ptrs), has_barrier=hasattr(h, "barrier"))
    t.fill_(float(rank))
    h.barrier()
    peer = h.get_buffer((rank + 1) % world, (16,), torch.float32)
    info["peer_value"] = float(peer[0].item())
    print("[symm]", info, flush=True)
except Exception as e:  # noqa: BLE001
    print("[symm] rank", rank, "FAILED:", type(e).__name__, str(e)[:500], flush=True)
can = [torch.cuda.can_device_access_peer(dev.index, j) for j in range(torch.cuda.device_count()) if j != dev.index]
print("[p2p] rank", rank, "can_access_peer:", can, flush=True)
dist.barrier()
dist.destroy_process_group()
