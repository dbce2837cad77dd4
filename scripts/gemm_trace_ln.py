"""Per-CTA phase timeline of the fused GEMM + LayerNorm launch against the plain GEMM (debug stamps), run on the GPU box."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf
from b200xf import kernels as K
from b200xf._lib import load, ptr

lib = load()
for (M, N, Kd) in [(4096, 512, 512), (4096, 512, 2048)]:
    x = torch.randn(M, Kd, device="cuda").to(torch.bfloat16); w = torch.randn(Kd, N, device="cuda").to(torch.bfloat16)
    r = torch.randn(M, N, device="cuda").to(torch.bfloat16)
    bias, gamma, beta = (torch.randn(N, device="cuda") for _ in range(3))
    for name, fn in (("plain", lambda: K.gemm(x, w, b_mn=True, bias=bias, resid=r)), ("fused LN", lambda: K.dense_layernorm_fwd(x, w, bias, r, gamma, beta))):
        tr = torch.zeros(148 * 8, device="cuda", dtype=torch.int64)
        for _ in range(3): fn()
        torch.cuda.synchronize()
        lib.b200xf_debug_gemm_trace(ptr(tr)); fn(); torch.cuda.synchronize(); lib.b200xf_debug_gemm_trace(None)
        t = tr.cpu().view(148, 8); used = t[:, 0] > 0; t = t[used]; t0 = t[:, 0].min()
        rel = (t - t0).float() / 1e3
        names = ["start", "setup", "1st data", "last mma", "acc ready", "epi(pass1) done", "exit", "after cluster sync"]
        print(f"{name} M={M} N={N} K={Kd}: {int(used.sum())} CTAs; us since first CTA start (median / max over CTAs)")
        print("   " + "  ".join(f"{n}: {float(rel[1:, i].median()):.2f}/{float(rel[1:, i].max()):.2f}" for i, n in enumerate(names)))
