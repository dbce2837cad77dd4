"""Per-kernel timeline of one graph-replayed train step, from CUPTI activity records (torch.profiler).  Diagnosis only:
timings under a profiler are not bench numbers.  Run on the GPU box:  python scripts/step_profile.py [config] > gpurun_out/step_kernels.txt"""
import sys
from pathlib import Path
import torch
from torch.profiler import profile, ProfilerActivity
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf

name = sys.argv[1] if len(sys.argv) > 1 else "base"
cfg = dict(b200xf.CONFIGS[name])
B, S, T = cfg["B"], cfg["S"], cfg["T"]
hp = dict(layers_count=cfg["layers_count"], embedding_size=cfg["embedding_size"], heads_count=cfg["heads_count"],
          feed_forward_hidden_size=cfg["feed_forward_hidden_size"], input_vocabulary_size=cfg["input_vocabulary_size"],
          target_vocabulary_size=cfg["target_vocabulary_size"], maximum_input_length=max(1000, S), maximum_target_length=max(1000, T + 1),
          droput_rate=0.0)
import os
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:  # torchrun: profile rank 0 of a data-parallel job
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
model = b200xf.Transformer(**hp, max_tokens=B * max(S, T + 1), seed=0)
model.trainable_variables
b200xf.B200Strategy(rank, world).attach(model)
model.iterations = 1
dev = torch.device("cuda", local)
inp, tgt = b200xf.synthetic_batch(B, S, T + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=1, ragged=False)
inp, tgt = torch.from_numpy(inp).to(dev), torch.from_numpy(tgt).to(dev)
for _ in range(8):
    model.train_step_device(inp, tgt, workers_count=world)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        model.train_step_device(inp, tgt, workers_count=world)
    torch.cuda.synchronize()
if rank != 0:
    model.close(); sys.exit(0)
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
if not evs:
    print("no CUDA activity records"); sys.exit(0)
# keep the last step: split at the largest gaps between prep_targets kernels
starts = [i for i, e in enumerate(evs) if "prep_targets" in e.name]
lo = starts[-1] if starts else 0
evs = evs[lo:]
t0 = evs[0].time_range.start
streams = {}
print(f"# {name}: {len(evs)} device activities in the last profiled step; columns: start_us dur_us end_us stream name")
for e in evs:
    sid = getattr(e, "device_index", 0)
    st = e.time_range.start - t0
    du = e.time_range.end - e.time_range.start
    nm = e.name.replace("b200xf::", "").replace("(anonymous namespace)::", "")
    print(f"{st:9.1f} {du:8.1f} {st + du:9.1f}  {nm[:90]}")
model.close()
