"""Development aid: does torch's symmetric memory give peer / multicast pointers on this box?
torchrun --nproc-per-node N scripts/symm_probe.py"""
import os
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty((1 << 20,), dtype=torch.float32, device=torch.device("cuda", local))
    h = symm.rendezvous(t, dist.group.WORLD.group_name)
    info = {"rank": rank, "buffer_ptrs": [hex(p) for p in h.buffer_ptrs], "multicast_ptr": hex(h.multicast_ptr) if h.multicast_ptr else None,
            "world": h.world_size}
    print(info, flush=True)
    # peer write test through a torch view of the peer buffer
    peer = (rank + 1) % world
    buf = h.get_buffer(peer, (16,), torch.float32)
    h.barrier()
    buf.fill_(float(rank + 1))
    h.barrier()
    print("rank", rank, "own buffer now", t[:4].tolist(), flush=True)
except Exception as exc:  # noqa: BLE001
    import traceback
    traceback.print_exc()
    print("rank", rank, "symmetric memory unavailable:", exc, flush=True)
print("p2p", [torch.cuda.can_device_access_peer(local, d) for d in range(torch.cuda.device_count()) if d != local])
dist.barrier()
dist.destroy_process_group()
