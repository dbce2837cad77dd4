"""Per-CTA phase timeline of one GEMM launch (debug stamps), run on the GPU box."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf
from b200xf import kernels as K
from b200xf._lib import load, ptr

LAY = {"fwd": (False, True), "dgrad": (False, False), "wgrad": (True, True)}
lib = load()
for (M, N, Kd, lay) in [(4096, 512, 512, "fwd"), (4096, 1536, 512, "fwd"), (4096, 2048, 512, "fwd"), (512, 512, 4096, "wgrad"), (4096, 512, 2048, "dgrad")]:
    a_mn, b_mn = LAY[lay]
    a = torch.randn((Kd, M) if a_mn else (M, Kd), device="cuda").to(torch.bfloat16)
    b = torch.randn((Kd, N) if b_mn else (N, Kd), device="cuda").to(torch.bfloat16)
    out = torch.zeros((M, N), device="cuda", dtype=torch.float32 if lay == "wgrad" else torch.bfloat16)
    tr = torch.zeros(148 * 8, device="cuda", dtype=torch.int64)
    for _ in range(3):
        K.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, accumulate=(lay == "wgrad"))
    torch.cuda.synchronize()
    lib.b200xf_debug_gemm_trace(ptr(tr))
    K.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, accumulate=(lay == "wgrad"))
    torch.cuda.synchronize()
    lib.b200xf_debug_gemm_trace(None)
    t = tr.cpu().view(148, 8)
    used = t[:, 0] > 0
    t = t[used]
    t0 = t[:, 0].min()
    rel = (t - t0).float() / 1e3
    names = ["start", "setup", "1st data", "last mma", "acc ready", "epi done", "exit"]
    print(f"{lay} M={M} N={N} K={Kd}: {int(used.sum())} CTAs; us since first CTA start (median / max over CTAs)")
    print("   " + "  ".join(f"{n}: {float(rel[:, i].median()):.2f}/{float(rel[:, i].max()):.2f}" for i, n in enumerate(names)))
