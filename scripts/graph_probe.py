"""Does replaying the whole train step as one CUDA graph beat stream launches?  (experiment)"""
import sys, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf

cfg = dict(b200xf.CONFIGS["base"]); B, S, T = cfg.pop("B"), cfg.pop("S"), cfg.pop("T")
model = b200xf.Transformer(**cfg, maximum_input_length=1000, maximum_target_length=1000, droput_rate=0.0, max_tokens=B * (T + 1))
model.trainable_variables; model.iterations = 1
inp, tgt = b200xf.synthetic_batch(B, S, T + 1, 32000, 32000, seed=1)
inp = torch.tensor(inp).cuda(); tgt = torch.tensor(tgt).cuda()
for _ in range(5): out = model.train_step_device(inp, tgt, 1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
t0 = time.perf_counter()
for _ in range(30): out = model.train_step_device(inp, tgt, 1)
t_enq = (time.perf_counter() - t0) / 30 * 1e3
e1.record(); torch.cuda.synchronize()
print(f"stream launches: {e0.elapsed_time(e1) / 30:.3f} ms/step (host enqueue {t_enq:.3f} ms/step), loss {float(out[0]):.4f}")
# capture on the stream the context was created with is not possible (legacy current stream); build a model on a side stream
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    m2 = b200xf.Transformer(**cfg, maximum_input_length=1000, maximum_target_length=1000, droput_rate=0.0, max_tokens=B * (T + 1))
    m2.trainable_variables; m2.iterations = 1
    for _ in range(3): out = m2.train_step_device(inp, tgt, 1)
    side.synchronize()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g, stream=side):
            out = m2.train_step_device(inp, tgt, 1)
        g.replay(); side.synchronize()
        e0.record(side)
        for _ in range(30): g.replay()
        e1.record(side); side.synchronize()
        print(f"graph replay:    {e0.elapsed_time(e1) / 30:.3f} ms/step, loss {float(out[0]):.4f}, iterations {m2.iterations}")
    except Exception as ex:
        print("capture failed:", repr(ex)[:400])
