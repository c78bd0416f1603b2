"""Probe: does torch symmetric memory give peer pointers and an NVLS multicast pointer on this box?"""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank = int(os.environ["RANK"]); torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
t = symm.empty(64 * 1024 * 1024, dtype=torch.uint8, device="cuda")
h = symm.rendezvous(t, dist.group.WORLD.group_name if hasattr(dist.group.WORLD, "group_name") else dist.group.WORLD)
if rank == 0:
    print("buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "multicast_ptr", hex(h.multicast_ptr) if h.multicast_ptr else None, "signal_pad_ptrs", len(h.signal_pad_ptrs), flush=True)
dist.barrier(); dist.destroy_process_group()
