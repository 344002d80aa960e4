"""Probe: does torch symmetric memory (multicast / NVLS) come up on this box?  torchrun --nproc-per-node 2"""
import os, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1 << 20, dtype=torch.float32, device=torch.device("cuda", rank))
    h = symm.rendezvous(t, dist.group.WORLD.group_name)
    print(rank, "rendezvous ok; buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "multicast_ptr", hex(h.multicast_ptr or 0),
          "signal_pad_ptrs", len(h.signal_pad_ptrs), flush=True)
except Exception as e:
    print(rank, "symm mem failed:", repr(e)[:500], flush=True)
dist.barrier(); dist.destroy_process_group()
