"""Warm, back-to-back timings of the non-GEMM kernels at Transformer-base shapes (CUDA-graph replay, CUDA events)."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf
from b200xf import kernels as K


def timeit(fn, reps=20):
    if "--once" in sys.argv:  # for `ncu`: every kernel twice (cold, warm), no graph, no timing
        fn(); fn(); torch.cuda.synchronize()
        return float("nan")
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
    B, S, H, d, dff, V = 32, 128, 8, 512, 2048, 32000
    if "big" in sys.argv: B, S, H, d, dff = 16, 256, 16, 1024, 4096
    if "long" in sys.argv: B, S = 8, 1024
    N = B * S
    dev = "cuda"
    bf = lambda *s: torch.randn(*s, device=dev).to(torch.bfloat16)
    x, r = bf(N, d), bf(N, d)
    gamma, beta = torch.randn(d, device=dev), torch.randn(d, device=dev)
    y, z, mean, rstd = K.layernorm_residual_fwd(x, None, gamma, beta)
    print(f"ln_fwd            [{N}x{d}]: {timeit(lambda: K.layernorm_residual_fwd(x, None, gamma, beta)):7.2f} us  (ideal {N*d*4/6.5e6:.2f})")
    scr = K.layernorm_bwd_scratch(N, d, dev)
    print(f"ln_bwd+reduce     [{N}x{d}]: {timeit(lambda: K.layernorm_residual_bwd(x, r, mean, rstd, gamma, scr)):7.2f} us  (ideal {N*d*6/6.5e6:.2f})")
    toks = torch.randint(4, 100, (B, S), device=dev)
    qkv = bf(B, S, 3 * d)
    q, k, v = qkv[:, :, :d], qkv[:, :, d:2 * d], qkv[:, :, 2 * d:]
    for causal in (False, True):
        o, lse = K.attention_fwd(q, k, v, toks, causal, H)
        fl = 4.0 * B * H * S * S * (d // H)
        t = timeit(lambda: K.attention_fwd(q, k, v, toks, causal, H))
        print(f"attn_fwd causal={int(causal)} B{B} H{H} S{S} dk{d//H}: {t:7.2f} us  {fl/t/1e6:6.1f} TF (dense count)")
        do = bf(B, S, d)
        t = timeit(lambda: K.attention_bwd(q, k, v, o, lse, do, toks, causal, H))
        print(f"attn_bwd causal={int(causal)}                  : {t:7.2f} us  {2.5*fl/t/1e6:6.1f} TF (dense count, 2 kernels)")
    lg = bf(N, V); labels = torch.randint(1, V, (N,), device=dev)
    print(f"vocab_ce (+count,+finalize) [{N}x{V}]: {timeit(lambda: K.vocab_ce_fwd_bwd(lg, labels, V, 1, True), reps=5):7.2f} us  (ideal {N*V*4/6.5e6:.1f})")
    n = 93_000_000 // 4 * 4
    w = torch.randn(n, device=dev); g = torch.randn(n, device=dev) * 1e-3; m = torch.zeros(n, device=dev); vv = torch.zeros(n, device=dev)
    sh = torch.empty(n, device=dev, dtype=torch.bfloat16); it = torch.tensor([10], device=dev)
    print(f"adam [{n}]: {timeit(lambda: K.adam_step(w, g, m, vv, it, d, w_bf16=sh, zero_grad=True), reps=3):7.2f} us  (ideal {n*34/6.5e6:.1f})")
    tab = torch.randn(V, d, device=dev); pe = K.positional_encoding(1000, d).to(dev)
    print(f"embed_fwd: {timeit(lambda: K.embed_pe_fwd(toks, tab, pe)):7.2f} us")


if __name__ == "__main__":
    main()
