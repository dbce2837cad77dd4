"""Timeline of a chain of back-to-back GEMM launches (debug stamps, globaltimer), run on the GPU box."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf
from b200xf import kernels as K
from b200xf._lib import load, ptr

LAY = {"fwd": (False, True), "dgrad": (False, False), "wgrad": (True, True)}
lib = load()
NCHAIN = 5
shapes = [(4096, 512, 512, "fwd"), (4096, 2048, 512, "fwd"), (4096, 512, 2048, "fwd"), (512, 512, 4096, "wgrad")]
if len(sys.argv) > 1:
    shapes = [tuple(int(x) if x.isdigit() else x for x in s.split(",")) for s in sys.argv[1:]]
for (M, N, Kd, lay) in shapes:
    a_mn, b_mn = LAY[lay]
    a = torch.randn((Kd, M) if a_mn else (M, Kd), device="cuda").to(torch.bfloat16)
    b = torch.randn((Kd, N) if b_mn else (N, Kd), device="cuda").to(torch.bfloat16)
    out = torch.zeros((M, N), device="cuda", dtype=torch.float32 if lay == "wgrad" else torch.bfloat16)
    trs = [torch.zeros(148 * 8, device="cuda", dtype=torch.int64) for _ in range(NCHAIN)]
    run = lambda: K.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, accumulate=(lay == "wgrad"))
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    for tr in trs:
        lib.b200xf_debug_gemm_trace(ptr(tr))
        run()
    lib.b200xf_debug_gemm_trace(None)
    torch.cuda.synchronize()
    ts = [tr.cpu().view(148, 8) for tr in trs]
    t0 = int(ts[0][:, 0][ts[0][:, 0] > 0].min())
    names = ["start", "setup", "1st data", "last mma", "acc ready", "epi done", "exit"]
    print(f"{lay} M={M} N={N} K={Kd}: us since the first CTA of launch 0 started; min / median / max over CTAs")
    for i, t in enumerate(ts):
        t = t[t[:, 0] > 0]
        rel = (t - t0).float() / 1e3
        print(f"  launch {i} ({t.shape[0]} CTAs): " + "  ".join(
            f"{n} {float(rel[:, j].min()):.2f}/{float(rel[:, j].median()):.2f}/{float(rel[:, j].max()):.2f}" for j, n in enumerate(names)))
