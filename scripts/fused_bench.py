"""Isolated timings of the fused GEMM epilogues against the kernels they replace (CUDA-graph replay of 20 back-to-back launches,
CUDA events).  python scripts/fused_bench.py [big]"""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf
from b200xf import kernels as K


def timeit(fn, reps=20):
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        fn(); side.synchronize()
        with torch.cuda.graph(g, stream=side):
            for _ in range(reps): fn()
        g.replay(); side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side); g.replay(); g.replay(); e1.record(side); side.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (2 * reps)


def main():
    N, d, dff, V = 4096, 512, 2048, 32000
    if "big" in sys.argv: d, dff = 1024, 4096
    dev = "cuda"
    bf = lambda *s: torch.randn(*s, device=dev).to(torch.bfloat16)
    gamma, beta, bias = torch.randn(d, device=dev), torch.randn(d, device=dev), torch.randn(d, device=dev)
    for kin in (d, dff):
        x, w, r = bf(N, kin), bf(kin, d), bf(N, d)
        z = torch.empty(N, d, device=dev, dtype=torch.bfloat16)
        t_g = timeit(lambda: K.gemm(x, w, b_mn=True, bias=bias, resid=r, out=z))
        t_l = timeit(lambda: K.layernorm_residual_fwd(z, None, gamma, beta))
        t_gl = timeit(lambda: (K.gemm(x, w, b_mn=True, bias=bias, resid=r, out=z), K.layernorm_residual_fwd(z, None, gamma, beta)))
        t_f = timeit(lambda: K.dense_layernorm_fwd(x, w, bias, r, gamma, beta))
        print(f"fwd [{N}x{kin}]x[{kin}x{d}] + LN: gemm {t_g:6.2f}  ln {t_l:6.2f}  gemm->ln {t_gl:6.2f}  fused {t_f:6.2f} us")
    for kin, nout in ((d, 3 * d), (d, d), (d, dff), (dff, d), (d, V)):
        x, dy = bf(N, kin), bf(N, nout)
        dw = torch.zeros(kin, nout, device=dev)
        t_w = timeit(lambda: K.gemm(x, dy, a_mn=True, b_mn=True, out=dw, accumulate=True))
        t_c = timeit(lambda: K.colsum(dy)) if hasattr(K, "colsum") else float("nan")
        t_f = timeit(lambda: K.dense_wgrad_bias(x, dy, dw))
        print(f"wgrad [{kin}x{N}]x[{N}x{nout}]: gemm {t_w:6.2f}  colsum kernel {t_c:6.2f}  fused (+reduce) {t_f:6.2f} us")


try:
    main()
except BrokenPipeError:
    pass
