"""Times the tcgen05 GEMM on the train step's shapes (CUDA-graph replay, CUDA events).  Run on the GPU box."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf
from b200xf import kernels as K

LAY = {"fwd": (False, True), "dgrad": (False, False), "wgrad": (True, True)}


def bench(M, N, Kd, lay, bn=0, split_k=0, reps=20):
    a_mn, b_mn = LAY[lay]
    pad = lambda n: (n + 7) // 8 * 8
    a = torch.randn((Kd, pad(M)) if a_mn else (M, pad(Kd)), device="cuda").to(torch.bfloat16)
    b = torch.randn((Kd, pad(N)) if b_mn else (N, pad(Kd)), device="cuda").to(torch.bfloat16)
    a = a[:, :M] if a_mn else a[:, :Kd]; b = b[:, :N] if b_mn else b[:, :Kd]
    out = torch.zeros((M, pad(N)), device="cuda", dtype=torch.float32 if lay == "wgrad" else torch.bfloat16)[:, :N]
    run = lambda: K.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, accumulate=(lay == "wgrad"), force_bn=bn, split_k=split_k)
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        run(); side.synchronize()
        with torch.cuda.graph(g, stream=side):
            for _ in range(reps): run()
        g.replay(); side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side); g.replay(); g.replay(); e1.record(side); side.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (2 * reps)
    return us, 2.0 * M * N * Kd / us / 1e6


if __name__ == "__main__":
    d, dff, V, tok = 512, 2048, 32000, 4096
    cases = [(tok, 3 * d, d, "fwd"), (tok, d, d, "fwd"), (tok, dff, d, "fwd"), (tok, d, dff, "fwd"), (tok, 12 * d, d, "fwd"), (tok, V, d, "fwd"),
             (tok, d, 3 * d, "dgrad"), (tok, d, d, "dgrad"), (tok, d, dff, "dgrad"), (tok, dff, d, "dgrad"), (tok, d, V, "dgrad"),
             (d, 3 * d, tok, "wgrad"), (d, d, tok, "wgrad"), (d, dff, tok, "wgrad"), (dff, d, tok, "wgrad"), (d, V, tok, "wgrad"),
             (8192, 8192, 8192, "dgrad"), (8192, 8192, 8192, "fwd")]
    lib = b200xf._lib.load()
    if "--pairs" in sys.argv:  # single-CTA kernel vs CTA pairs (cta_group::2) vs the heuristic
        for M, N, Kd, lay in cases:
            cols = []
            for mode, bn, tag in ((0, 0, "single auto"), (1, 128, "pair bn128"), (1, 256, "pair bn256"), (-1, 0, "heuristic")):
                lib.b200xf_gemm_set_pair_mode(mode)
                us, tf = bench(M, N, Kd, lay, bn)
                cols.append(f"{tag}: {us:7.1f} us {tf:7.1f} TF")
            lib.b200xf_gemm_set_pair_mode(-1)
            print(f"{lay:6s} M={M:5d} N={N:5d} K={Kd:5d} | " + " | ".join(cols), flush=True)
        sys.exit(0)
    for M, N, Kd, lay in cases:
        res = []
        for bn in (64, 128, 256):
            us, tf = bench(M, N, Kd, lay, bn)
            res.append(f"bn{bn}: {us:7.1f} us {tf:7.1f} TF")
        us, tf = bench(M, N, Kd, lay, 0)
        print(f"{lay:6s} M={M:5d} N={N:5d} K={Kd:5d} | " + " | ".join(res) + f" | auto: {us:7.1f} us {tf:7.1f} TF", flush=True)
