"""Launches the step's dominant GEMM (vocab projection forward, M=4096 N=32000 K=512) a few times: target for `ncu --set full`."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf
from b200xf import kernels as K

M, N, Kd = 4096, 32000, 512
x = torch.randn(M, Kd, device="cuda").to(torch.bfloat16)
w = torch.randn(Kd, N, device="cuda").to(torch.bfloat16)
bias = torch.randn(N, device="cuda")
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(6):
    K.gemm(x, w, b_mn=True, bias=bias, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    K.gemm(x, w, b_mn=True, bias=bias, out=out)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 100
print(f"vocab fwd GEMM: {us:.1f} us, {2.0 * M * N * Kd / us / 1e6:.1f} TFLOP/s")
