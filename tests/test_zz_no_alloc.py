"""GPU test of b200xf_layernorm_residual_bwd_noalloc, the LayerNorm backward for callers that cannot allocate device
memory (an op inside a framework whose pool already owns the GPU, a loop that must avoid cudaMalloc's device sync).

It is called through the C ABI twice in a row on different problems — the second smaller than the first, so the
library's built-in scratch array still holds the first call's partial sums — and each result is compared with the
ordinary entry point (caller-provided scratch) on the same inputs.  Both run the same kernels in the same order, so the
comparison uses the tight end of this suite's LayerNorm tolerances.  Sizes past the built-in array are refused.
This file sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

import b200xf
from b200xf import kernels as K
from b200xf._lib import load, ptr, stream_ptr

DEV = "cuda"


def _noalloc_bwd(dy, z, mean, rstd, gamma):
    rows, d = dy.shape
    dz = torch.empty_like(dy)
    dgamma = torch.empty(d, device=DEV); dbeta = torch.empty(d, device=DEV); dcol = torch.empty(d, device=DEV)
    rc = load().b200xf_layernorm_residual_bwd_noalloc(ptr(dy), ptr(z), ptr(mean), ptr(rstd), ptr(gamma), ptr(dz), ptr(dgamma),
                                                      ptr(dbeta), ptr(dcol), rows, d, stream_ptr())
    return rc, (dz, dgamma, dbeta, dcol)


def test_noalloc_layernorm_bwd_matches_scratch_variant_twice_in_a_row():
    g = torch.Generator(device=DEV).manual_seed(23)
    for rows, d in [(4096, 512), (300, 128)]:  # two calls in a row; the second leaves most of the first one's partials in place
        x = torch.randn(rows, d, device=DEV, generator=g).to(torch.bfloat16)
        r = torch.randn(rows, d, device=DEV, generator=g).to(torch.bfloat16)
        gamma = torch.randn(d, device=DEV, generator=g); beta = torch.randn(d, device=DEV, generator=g)
        dy = torch.randn(rows, d, device=DEV, generator=g).to(torch.bfloat16)
        _, z, mean, rstd = K.layernorm_residual_fwd(x, r, gamma, beta, 1e-6, want_z=True)
        want = K.layernorm_residual_bwd(dy, z, mean, rstd, gamma)
        rc, got = _noalloc_bwd(dy, z, mean, rstd, gamma)
        assert rc == 0, load().b200xf_last_error()
        torch.cuda.synchronize()
        for name, a, b in zip(("dz", "dgamma", "dbeta", "dcol"), got, want):
            assert torch.isfinite(a.float()).all(), name
            err = float((a.float() - b.float()).norm() / (b.float().norm() + 1e-30))
            assert err < 1e-5, (rows, d, name, err)  # fp32-sum tolerance of test_layernorm_fwd_bwd (dcol: 1e-5)
        # and directly against torch's fp32 LayerNorm backward on the same saved z (independent of the sibling entry point)
        zf = z.float().requires_grad_(True); gf = gamma.clone().requires_grad_(True); bf = beta.clone().requires_grad_(True)
        torch.nn.functional.layer_norm(zf, (d,), gf, bf, 1e-6).backward(dy.float())
        for name, a, b, tol in (("dz", got[0], zf.grad, 4e-3), ("dgamma", got[1], gf.grad, 2e-4), ("dbeta", got[2], bf.grad, 2e-4)):
            err = float((a.float() - b).norm() / (b.norm() + 1e-30))
            assert err < tol, (rows, d, name, err)  # dz is bf16 (one rounding); the sums are fp32 (test_layernorm_fwd_bwd's bounds)


@pytest.mark.parametrize("d,fits,too_big", [(512, 9472, 9488), (1024, 2368, 2376)])
def test_noalloc_layernorm_bwd_capacity_edge(d, fits, too_big):
    """The largest problem the built-in array holds is accepted and correct; one block of rows more is refused."""
    g = torch.Generator(device=DEV).manual_seed(5)
    gamma = torch.randn(d, device=DEV, generator=g); beta = torch.randn(d, device=DEV, generator=g)
    x = torch.randn(too_big, d, device=DEV, generator=g).to(torch.bfloat16)
    r = torch.randn(too_big, d, device=DEV, generator=g).to(torch.bfloat16)
    dy = torch.randn(too_big, d, device=DEV, generator=g).to(torch.bfloat16)
    _, z, mean, rstd = K.layernorm_residual_fwd(x, r, gamma, beta, 1e-6, want_z=True)
    rc, got = _noalloc_bwd(dy[:fits], z[:fits], mean[:fits], rstd[:fits], gamma)
    assert rc == 0, load().b200xf_last_error()
    want = K.layernorm_residual_bwd(dy[:fits].contiguous(), z[:fits].contiguous(), mean[:fits].contiguous(), rstd[:fits].contiguous(), gamma)
    torch.cuda.synchronize()
    for name, a, b in zip(("dz", "dgamma", "dbeta", "dcol"), got, want):
        err = float((a.float() - b.float()).norm() / (b.float().norm() + 1e-30))
        assert err < 1e-5, (d, name, err)
    rc, _ = _noalloc_bwd(dy, z, mean, rstd, gamma)
    assert rc != 0 and "built-in buffer" in load().b200xf_last_error().decode()


def test_noalloc_layernorm_bwd_rejects_sizes_beyond_its_buffer():
    rows, d = 4096, 2048  # 512 blocks x 3 x 2048 floats = 12.6 MB > the 3.6 MB built-in array
    t = torch.zeros(rows, d, device=DEV, dtype=torch.bfloat16)
    v = torch.zeros(rows, device=DEV); gm = torch.ones(d, device=DEV)
    rc, _ = _noalloc_bwd(t, t, v, v, gm)
    assert rc != 0
    msg = load().b200xf_last_error().decode()
    assert "built-in buffer" in msg and "scratch" in msg
