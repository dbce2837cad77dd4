"""Bandwidth of the fused NVLS update kernel (reduce over replicas + Adam + broadcast) against NCCL's reduce-scatter + all-gather on
the same range.  torchrun --nproc-per-node N scripts/nvls_bench.py"""
import os, sys
from pathlib import Path
import torch
import torch.distributed as dist
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf
from b200xf._lib import load, check, ptr

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cfg = dict(b200xf.CONFIGS["base"])
hp = {k: cfg[k] for k in ("layers_count", "embedding_size", "heads_count", "feed_forward_hidden_size", "input_vocabulary_size", "target_vocabulary_size")}
model = b200xf.Transformer(**hp, maximum_input_length=1000, maximum_target_length=1000, droput_rate=0.0, max_tokens=4096 + 64)
model.trainable_variables
b200xf.B200Strategy(rank, world).attach(model)
model.iterations = 10
lib = load()
assert model._symmetric is not None, "NVLS not attached"
for mb in ((32, 64) if world > 2 else (8, 32, 64, 128)):
    n = (mb << 18) // (world * 64) * world * 64
    for ctas in ((32, 64, 148) if world > 2 else (16, 32, 64, 148, 296)):
        for refresh in (0,):
            for _ in range(3): check(lib.b200xf_debug_nvls_update(model._ctx, 0, n, ctas, refresh))
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10): check(lib.b200xf_debug_nvls_update(model._ctx, 0, n, ctas, refresh))
            e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 100
            if rank == 0:
                print(f"nvls  {mb:4d} MB  ctas {ctas:3d}: {us:8.1f} us   algbw {n * 4 / us / 1e3:7.1f} GB/s  (shard {n * 4 / world / 1e6:.1f} MB in + out per rank)", flush=True)
    buf = torch.zeros(n, device="cuda")
    for _ in range(3): check(lib.b200xf_allreduce_bucket(model._ctx, ptr(buf), n))
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): check(lib.b200xf_allreduce_bucket(model._ctx, ptr(buf), n))
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 100
    if rank == 0:
        print(f"nccl allreduce {mb:4d} MB: {us:8.1f} us   algbw {n * 4 / us / 1e3:7.1f} GB/s", flush=True)
assert lib.b200xf_nvls_error(model._ctx) == 0
model.close()
dist.destroy_process_group()
