"""Probe: torch symmetric memory rendezvous + NVLS multicast availability (run under torchrun, >= 2 GPUs)."""
import os, sys, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{lr}"))
import torch.distributed._symmetric_memory as sm
t = sm.empty(1 << 20, dtype=torch.uint8, device=f"cuda:{lr}")
t.zero_()
h = sm.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "mc", hex(h.multicast_ptr), "pads", len(h.signal_pad_ptrs), flush=True)
print(rank, [a for a in dir(h) if not a.startswith("_")], flush=True)
dist.barrier(); dist.destroy_process_group()
