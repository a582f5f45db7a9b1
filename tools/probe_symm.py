"""Dev (GPU, torchrun): is torch symmetric memory with an NVLS multicast pointer available on this box?"""
import os

import torch
import torch.distributed as dist

rank = int(os.environ["RANK"])
world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
try:
    import torch.distributed._symmetric_memory as symm

    t = symm.empty(1 << 20, dtype=torch.uint8, device="cuda")
    h = symm.rendezvous(t, dist.group.WORLD.group_name)
    info = dict(rank=rank, world=h.world_size, has_mc=bool(getattr(h, "multicast_ptr", 0)),
                mc_ptr=hex(getattr(h, "multicast_ptr", 0)), buffer_ptrs=[hex(p) for p in h.buffer_ptrs][:4],
                signal_pad=[hex(p) for p in h.signal_pad_ptrs][:2])
    print(info, flush=True)
    # peer write visibility test through the peer pointers
    t.fill_(rank + 1)
    h.barrier()
    peer = h.get_buffer((rank + 1) % world, (16,), torch.uint8)
    print(rank, "peer value", int(peer[0]), flush=True)
    h.barrier()
except Exception as e:  # noqa: BLE001
    import traceback

    traceback.print_exc()
    print(rank, "symmetric memory unavailable:", type(e).__name__, e, flush=True)
dist.destroy_process_group()
