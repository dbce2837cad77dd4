"""Does this box give torch symmetric memory a multicast (NVLS) mapping?  torchrun --nproc-per-node N scripts/symm_probe.py"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
t = symm_mem.empty(1 << 20, dtype=torch.float32, device=f"cuda:{local}")
h = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
print(f"rank {rank}: world {h.world_size} multicast_ptr {h.multicast_ptr:#x} buffer_ptrs {[hex(p) for p in h.buffer_ptrs]} "
      f"signal_pad_ptrs {[hex(p) for p in h.signal_pad_ptrs]} signal_pad_size {h.signal_pad_size}", flush=True)
t.fill_(rank + 1.0)
h.barrier(0)
peer = h.get_buffer((rank + 1) % h.world_size, (8,), torch.float32)
print(f"rank {rank}: peer buffer head {peer.tolist()}", flush=True)
h.barrier(0)
dist.destroy_process_group()
