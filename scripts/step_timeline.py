"""Timeline of the tcgen05 GEMM launches inside one graph-replayed train step (debug stamps, %globaltimer).
The gaps between GEMMs are the other kernels (LayerNorm, attention, loss, ...) plus idle time.  Run on the GPU box:
    python scripts/step_timeline.py [config] > gpurun_out/timeline.txt
"""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf
from b200xf._lib import load, ptr

name = sys.argv[1] if len(sys.argv) > 1 else "base"
cfg = dict(b200xf.CONFIGS[name])
B, S, T = cfg["B"], cfg["S"], cfg["T"]
hp = dict(layers_count=cfg["layers_count"], embedding_size=cfg["embedding_size"], heads_count=cfg["heads_count"],
          feed_forward_hidden_size=cfg["feed_forward_hidden_size"], input_vocabulary_size=cfg["input_vocabulary_size"],
          target_vocabulary_size=cfg["target_vocabulary_size"], maximum_input_length=max(1000, S), maximum_target_length=max(1000, T + 1),
          droput_rate=0.0)
model = b200xf.Transformer(**hp, max_tokens=B * max(S, T + 1), seed=0)
model.trainable_variables  # build
model.iterations = 1
dev = torch.device("cuda:0")
inp, tgt = b200xf.synthetic_batch(B, S, T + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=1, ragged=False)
inp, tgt = torch.from_numpy(inp).to(dev), torch.from_numpy(tgt).to(dev)
lib = load()
SLOTS = 256
ring = torch.zeros(SLOTS * 148 * 8, device=dev, dtype=torch.int64)
lib.b200xf_debug_gemm_trace_ring(ptr(ring), SLOTS)   # before the graph is captured: the slots are baked into it
for _ in range(8):
    model.train_step_device(inp, tgt)
torch.cuda.synchronize()
ring.zero_()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); model.train_step_device(inp, tgt); e1.record()
torch.cuda.synchronize()
lib.b200xf_debug_gemm_trace_ring(None, 0)
t = ring.cpu().view(SLOTS, 148, 8)
rows = []
for s in range(SLOTS):
    used = t[s, :, 0] > 0
    if not bool(used.any()):
        continue
    u = t[s][used]
    d = int(t[s, 0, 7])
    M, N, K, a_mn, b_mn = d >> 44, (d >> 24) & 0xFFFFF, (d >> 4) & 0xFFFFF, (d >> 1) & 1, d & 1
    kind = "wgrad" if (a_mn and b_mn) else ("dgrad" if (not a_mn and not b_mn) else "fwd")
    rows.append(dict(slot=s, start=int(u[:, 0].min()), ready=int(u[:, 1].median()), data=int(u[:, 2][u[:, 2] > 0].median()) if bool((u[:, 2] > 0).any()) else 0,
                     end=int(u[:, 6].max()), ctas=int(used.sum()), M=M, N=N, K=K, kind=kind))
rows.sort(key=lambda r: r["start"])
t0 = rows[0]["start"]
print(f"# {name}: step {e0.elapsed_time(e1):.3f} ms by CUDA events; {len(rows)} GEMM launches traced (tracing adds a little time)")
print("# columns: first CTA start | dependencies resolved (median CTA, after griddepcontrol.wait) | last CTA exit | busy = exit - resolved |"
      " gap = resolved - previous GEMM's exit on the same stream class (the kernels in between + idle) | kind M N K ctas; side-stream launches indented")
last_end = {"main": None, "side": None}
tot = {"main": 0.0, "side": 0.0}; gaps = {"main": 0.0, "side": 0.0}
for r in rows:
    cls = "side" if r["kind"] == "wgrad" else "main"
    st, rd, en = (r["start"] - t0) / 1e3, (r["ready"] - t0) / 1e3, (r["end"] - t0) / 1e3
    gap = rd - last_end[cls] if last_end[cls] is not None else 0.0
    last_end[cls] = en
    tot[cls] += en - rd; gaps[cls] += max(gap, 0.0)
    ind = "        " if cls == "side" else ""
    print(f"{ind}{st:9.2f} {rd:9.2f} {en:9.2f} {en - rd:7.2f} {gap:8.2f}  {r['kind']:5s} {r['M']:5d} {r['N']:5d} {r['K']:5d}  {r['ctas']:3d}")
span = (max(r["end"] for r in rows) - t0) / 1e3
print(f"# first GEMM start -> last GEMM end: {span:.1f} us; main-stream GEMM time {tot['main']:.1f} us, gaps between them {gaps['main']:.1f} us; "
      f"side-stream GEMM time {tot['side']:.1f} us")
model.close()
