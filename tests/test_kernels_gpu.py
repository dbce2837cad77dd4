"""Per-kernel parity on the GPU, through the C ABI (ctypes): each CUDA kernel against a plain torch fp32
restatement of the same op on the same (bf16-rounded) inputs.  Tolerances are stated per test."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import b200xf
from b200xf import kernels as K

DEV = "cuda"


def _bf(x):
    return x.to(torch.bfloat16)


def _rel(a, b):
    a = a.float(); b = b.float()
    return float((a - b).norm() / (b.norm() + 1e-30))


# ------------------------------------------------------------------------------------------ GEMM
LAYOUTS = {"fwd": (False, True), "dgrad": (False, False), "wgrad": (True, True), "mn_k": (True, False)}
SHAPES = [(128, 128, 64), (256, 256, 128), (300, 200, 136), (2560, 384, 128), (130, 7010, 128), (512, 520, 4096)]


def _make_operand(rows_logical, k, mn_major, g):
    """Returns (stored tensor view with 8-aligned pitch, logical [rows, k] fp32 of the bf16 values)."""
    if mn_major:  # stored [k, rows]
        ld = (rows_logical + 7) // 8 * 8
        buf = _bf(torch.randn(k, ld, device=DEV, generator=g))
        view = buf[:, :rows_logical]
        return view, view.float().t().contiguous()
    ld = (k + 7) // 8 * 8
    buf = _bf(torch.randn(rows_logical, ld, device=DEV, generator=g))
    view = buf[:, :k]
    return view, view.float()


@pytest.mark.parametrize("layout", list(LAYOUTS))
@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("bn", [0, 64, 128, 256])
def test_gemm_layouts(layout, shape, bn):
    M, N, Kd = shape
    a_mn, b_mn = LAYOUTS[layout]
    g = torch.Generator(device=DEV).manual_seed(hash((layout, shape)) % 2**31)
    a, a_log = _make_operand(M, Kd, a_mn, g)
    b, b_log = _make_operand(N, Kd, b_mn, g)
    ref = a_log @ b_log.t()
    ldn = (N + 7) // 8 * 8
    out = torch.full((M, ldn), float("nan"), device=DEV, dtype=torch.float32)[:, :N]
    K.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, force_bn=bn)
    torch.cuda.synchronize()
    assert torch.isfinite(out).all(), f"non-finite output {layout} {shape} bn={bn}"
    err = _rel(out, ref)
    assert err < 2e-5, f"{layout} {shape} bn={bn}: rel err {err}"  # fp32 accumulate of exact bf16 products


def test_gemm_epilogues():
    g = torch.Generator(device=DEV).manual_seed(7)
    M, N, Kd = 384, 320, 192
    x = _bf(torch.randn(M, Kd, device=DEV, generator=g)); w = _bf(torch.randn(Kd, N, device=DEV, generator=g) * 0.1)
    bias = torch.randn(N, device=DEV, generator=g); resid = _bf(torch.randn(M, N, device=DEV, generator=g))
    base = x.float() @ w.float()
    y = K.gemm(x, w, b_mn=True, bias=bias)
    assert _rel(y, base + bias) < 4e-3  # bf16 output rounding (2^-9 relative per element)
    y = K.gemm(x, w, b_mn=True, bias=bias, relu=True)
    assert _rel(y, torch.relu(base + bias)) < 4e-3
    y = K.gemm(x, w, b_mn=True, bias=bias, resid=resid, resid_op=1)
    assert _rel(y, base + bias + resid.float()) < 4e-3
    y = K.gemm(x, w, b_mn=True, resid=resid, resid_op=2)
    assert _rel(y, base * (resid.float() > 0)) < 4e-3
    y = K.gemm(x, w, b_mn=True, alpha=0.5, out_fp32=True)
    assert _rel(y, 0.5 * base) < 2e-5


@pytest.mark.parametrize("split_k", [0, 1, 3, 8])
def test_gemm_wgrad_accumulate_split_k(split_k):
    g = torch.Generator(device=DEV).manual_seed(11)
    tokens, Kd, N = 4096, 256, 200
    x = _bf(torch.randn(tokens, Kd, device=DEV, generator=g)); dy = _bf(torch.randn(tokens, 208, device=DEV, generator=g))[:, :N]
    acc = torch.ones((Kd, 208), device=DEV, dtype=torch.float32)[:, :N]
    K.gemm(x, dy, a_mn=True, b_mn=True, out=acc, accumulate=True, split_k=split_k)
    ref = 1.0 + x.float().t() @ dy.float()
    assert _rel(acc, ref) < 2e-5


@pytest.fixture
def forced_pairs():
    """Every eligible GEMM runs as CTA pairs (cta_group::2) inside the block; the heuristic is restored afterwards."""
    lib = b200xf._lib.load()
    b200xf._lib.check(lib.b200xf_gemm_set_pair_mode(1))
    yield
    b200xf._lib.check(lib.b200xf_gemm_set_pair_mode(-1))


@pytest.mark.parametrize("layout", list(LAYOUTS))
@pytest.mark.parametrize("shape", [(256, 256, 128), (300, 200, 136), (2560, 384, 128), (130, 7010, 128), (512, 520, 4096),
                                   (2176, 1000, 512), (1024, 4096, 96)])
@pytest.mark.parametrize("bn", [0, 128, 256])
def test_gemm_cta_pairs(layout, shape, bn, forced_pairs):
    # odd numbers of 128-row blocks (300, 130, 2176 rows: the last pair's second CTA is partly or wholly out of range),
    # ragged N and K, one and many tiles per pair
    test_gemm_layouts(layout, shape, bn)


def test_gemm_cta_pairs_epilogues_and_split_k(forced_pairs):
    test_gemm_epilogues()
    for split_k in (0, 1, 3, 8):
        test_gemm_wgrad_accumulate_split_k(split_k)


def test_gemm_cta_pairs_heuristic_matches_single():
    """The shapes the heuristic sends to CTA pairs (>= 3 waves of 256-wide tiles) agree with the single-CTA kernel."""
    lib = b200xf._lib.load()
    g = torch.Generator(device=DEV).manual_seed(5)
    x = _bf(torch.randn(4096, 512, device=DEV, generator=g)); w = _bf(torch.randn(512, 8192, device=DEV, generator=g) * 0.05)
    bias = torch.randn(8192, device=DEV, generator=g)
    try:
        b200xf._lib.check(lib.b200xf_gemm_set_pair_mode(0)); single = K.gemm(x, w, b_mn=True, bias=bias)
        b200xf._lib.check(lib.b200xf_gemm_set_pair_mode(-1)); auto = K.gemm(x, w, b_mn=True, bias=bias)
    finally:
        b200xf._lib.check(lib.b200xf_gemm_set_pair_mode(-1))
    torch.cuda.synchronize()
    assert torch.equal(single, auto)  # same products, same k order, same rounding


# ------------------------------------------------------------------------------------------ attention
def _attn_ref(q, k, v, key_tokens, causal, H):
    """transformer_model.py:73-109 + masks :56-70,:350-363, in fp32."""
    B, Sq, D = q.shape; Sk = k.shape[1]; dk = D // H
    sp = lambda x: x.float().reshape(B, -1, H, dk).permute(0, 2, 1, 3)
    logits = sp(q) @ sp(k).transpose(-1, -2) / math.sqrt(dk)
    mask = (key_tokens == 0).float()[:, None, None, :]
    if causal:
        la = 1.0 - torch.tril(torch.ones(Sq, Sk, device=q.device))
        mask = torch.maximum(mask, la)
    logits = logits + mask * -1e9
    w = torch.softmax(logits, -1)
    return (w @ sp(v)).permute(0, 2, 1, 3).reshape(B, Sq, D), w


ATTN_CASES = [  # B, H, Sq, Sk, dk, causal, pad_first_token
    (2, 8, 40, 40, 16, False, False), (2, 8, 40, 40, 16, True, False), (3, 4, 40, 33, 32, False, False),
    (2, 8, 128, 128, 64, True, False), (2, 8, 128, 128, 64, False, False), (1, 2, 200, 77, 64, False, False),
    (2, 2, 70, 70, 16, True, True), (1, 2, 300, 300, 64, True, False),
    # single-tile tcgen05 backward (Sq, Sk <= 128, dk 64): ragged tile, cross shapes, fully masked causal rows
    (3, 4, 100, 100, 64, True, False), (2, 8, 128, 77, 64, False, False), (2, 4, 40, 128, 64, False, False),
    (2, 2, 128, 128, 64, True, True), (5, 3, 17, 9, 64, False, False),
    # multi-tile tcgen05 backward (head size 64, longer than one tile): ragged tails, cross shapes, causal skipping, masked rows
    (2, 4, 256, 256, 64, True, False), (1, 2, 300, 300, 64, False, False), (2, 2, 200, 450, 64, False, False),
    (1, 2, 384, 130, 64, False, False), (2, 2, 257, 257, 64, True, True), (1, 8, 1024, 1024, 64, True, False),
]


@pytest.mark.parametrize("case", ATTN_CASES)
def test_attention_fwd_bwd(case):
    B, H, Sq, Sk, dk, causal, pad_first = case
    g = torch.Generator(device=DEV).manual_seed(3)
    D = H * dk
    if Sq == Sk:  # self-attention: Q|K|V are column slices of one fused buffer (row pitch 3D)
        qkv = _bf(torch.randn(B, Sq, 3 * D, device=DEV, generator=g))
        q = qkv[:, :, :D]; k = qkv[:, :, D:2 * D]; v = qkv[:, :, 2 * D:]
    else:         # cross-attention: Q [B,Sq,D], K|V column slices of a [B,Sk,2D] buffer
        q = _bf(torch.randn(B, Sq, D, device=DEV, generator=g))
        kv = _bf(torch.randn(B, Sk, 2 * D, device=DEV, generator=g))
        k = kv[:, :, :D]; v = kv[:, :, D:]
    toks = torch.randint(4, 100, (B, Sk), device=DEV, generator=g)
    for b in range(B):
        toks[b, Sk - (b * 5 + 3) % (Sk // 2):] = 0  # ragged right padding
    if pad_first:
        toks[0, :3] = 0  # fully masked causal rows -> must come out uniform like the reference
    o, lse = K.attention_fwd(q, k, v, toks, causal, H)
    qr, kr, vr = (t.float().detach().requires_grad_(True) for t in (q, k, v))
    ref, _ = _attn_ref(qr, kr, vr, toks, causal, H)
    assert torch.isfinite(o.float()).all()
    err = float((o.float() - ref.detach()).abs().max())
    assert err < 2e-2, f"{case}: fwd max abs err {err}"  # bf16 probabilities and output
    d_o = _bf(torch.randn(B, Sq, D, device=DEV, generator=g))
    ref.backward(d_o.float())
    for scratch in (True, False):  # with the fp32 dQ accumulator (tcgen05 paths) and without it (mma.sync path for long sequences)
        dq, dk_, dv = K.attention_bwd(q, k, v, o, lse, d_o, toks, causal, H, dq_scratch=scratch)
        for name, got, want in (("dq", dq, qr.grad), ("dk", dk_, kr.grad), ("dv", dv, vr.grad)):
            assert torch.isfinite(got.float()).all(), f"{case}: {name} not finite"
            e = _rel(got, want)
            assert e < 2e-2, f"{case} scratch={scratch}: {name} rel err {e}"


def test_attention_fwd_tcgen05_multi_tile_opt_in(monkeypatch):
    """The opt-in multi-tile tcgen05 forward (B200XF_ATTENTION_TC_FWD_MULTI=1) against the same reference."""
    monkeypatch.setenv("B200XF_ATTENTION_TC_FWD_MULTI", "1")
    g = torch.Generator(device=DEV).manual_seed(5)
    for (B, H, Sq, Sk, causal) in [(2, 2, 300, 300, True), (1, 4, 200, 450, False), (1, 2, 1024, 1024, True)]:
        D = H * 64
        q = _bf(torch.randn(B, Sq, D, device=DEV, generator=g)); kv = _bf(torch.randn(B, Sk, 2 * D, device=DEV, generator=g))
        k, v = kv[:, :, :D], kv[:, :, D:]
        toks = torch.randint(4, 100, (B, Sk), device=DEV, generator=g); toks[:, Sk - 37:] = 0
        o, lse = K.attention_fwd(q, k, v, toks, causal, H)
        ref, _ = _attn_ref(q, k, v, toks, causal, H)
        assert float((o.float() - ref).abs().max()) < 2e-2
        monkeypatch.setenv("B200XF_ATTENTION_TC_FWD_MULTI", "0")
        o2, lse2 = K.attention_fwd(q, k, v, toks, causal, H)
        monkeypatch.setenv("B200XF_ATTENTION_TC_FWD_MULTI", "1")
        assert float((o.float() - o2.float()).abs().max()) < 2e-2
        assert float((lse.sum(-1) - lse2.sum(-1)).abs().max()) < 2e-2     # max + log2(sum) agree


# ------------------------------------------------------------------------------------------ LayerNorm
@pytest.mark.parametrize("rows,d", [(2560, 128), (4096, 512), (1000, 1024), (37, 264)])
def test_layernorm_fwd_bwd(rows, d):
    g = torch.Generator(device=DEV).manual_seed(5)
    x = _bf(torch.randn(rows, d, device=DEV, generator=g)); r = _bf(torch.randn(rows, d, device=DEV, generator=g))
    gamma = torch.randn(d, device=DEV, generator=g); beta = torch.randn(d, device=DEV, generator=g)
    y, z, mean, rstd = K.layernorm_residual_fwd(x, r, gamma, beta, 1e-6, want_z=True)
    zr = _bf(x.float() + r.float()).float().requires_grad_(True)
    assert torch.equal(z.float(), zr.detach())
    gr = gamma.clone().requires_grad_(True); br = beta.clone().requires_grad_(True)
    ref = torch.nn.functional.layer_norm(zr, (d,), gr, br, 1e-6)
    assert float((y.float() - ref.detach()).abs().max()) < 4e-2 and _rel(y, ref.detach()) < 4e-3  # bf16 output
    np.testing.assert_allclose(mean.cpu(), zr.detach().mean(-1).cpu(), atol=1e-5)
    dy = _bf(torch.randn(rows, d, device=DEV, generator=g))
    y2, _, mean2, rstd2 = K.layernorm_residual_fwd(z, None, gamma, beta, 1e-6)
    assert torch.equal(y2, y)
    dz, dgamma, dbeta, dcol = K.layernorm_residual_bwd(dy, z, mean, rstd, gamma)
    ref.backward(dy.float())
    assert _rel(dz, zr.grad) < 4e-3
    assert _rel(dgamma, gr.grad) < 1e-4 and _rel(dbeta, br.grad) < 1e-4  # fp32 sums
    assert _rel(dcol, dz.float().sum(0)) < 1e-5
    # run twice with the same scratch: the counter must have been left at zero; result bit-identical (deterministic)
    scratch = K.layernorm_bwd_scratch(rows, d, DEV)
    a = K.layernorm_residual_bwd(dy, z, mean, rstd, gamma, scratch)
    b = K.layernorm_residual_bwd(dy, z, mean, rstd, gamma, scratch)
    for u, w in zip(a, b):
        assert torch.equal(u, w)


# ------------------------------------------------------------------------------------------ embedding
def test_embedding_fwd_bit_exact_and_bwd():
    g = torch.Generator(device=DEV).manual_seed(9)
    V, d, B, S = 7765, 128, 64, 40
    table = (torch.rand(V, d, device=DEV, generator=g) - 0.5) * 0.1
    pe = K.positional_encoding(1000, d).to(DEV)
    toks = torch.randint(0, V, (B, S), device=DEV, generator=g)
    toks[:, -5:] = 0
    out = K.embed_pe_fwd(toks, table, pe)
    ref = table[toks] * np.float32(math.sqrt(d)) + pe[:S]
    assert torch.equal(out, _bf(ref))  # gather bit-exact; one fp32 mul+add then a bf16 round
    dx = _bf(torch.randn(B, S, d, device=DEV, generator=g))
    dt = K.embed_pe_bwd(toks, dx, V)
    want = torch.zeros(V, d, device=DEV).index_add_(0, toks.reshape(-1), dx.float().reshape(-1, d)) * np.float32(math.sqrt(d))
    assert _rel(dt, want) < 1e-6
    assert torch.equal(dt, K.embed_pe_bwd(toks, dx, V))  # deterministic
    assert torch.count_nonzero(dt[torch.bincount(toks.reshape(-1), minlength=V) == 0]) == 0


def test_positional_encoding_matches_reference_kat():
    pe = K.positional_encoding(1000, 128)
    assert pe[0].tolist() == [0.0, 1.0] * 64
    np.testing.assert_allclose(pe[1, :4].numpy(), [0.841471, 0.5403023, 0.7617204, 0.6479058], rtol=2e-6)
    np.testing.assert_allclose(pe[1, -2:].numpy(), [1.1547821e-4, 1.0], rtol=2e-6)


# ------------------------------------------------------------------------------------------ cross-entropy
@pytest.mark.parametrize("rows,V,workers", [(2560, 7010, 1), (512, 32000, 3), (64, 50, 1), (128, 12345, 1), (64, 40001, 2)])
def test_vocab_ce(rows, V, workers):
    g = torch.Generator(device=DEV).manual_seed(13)
    ld = (V + 7) // 8 * 8
    buf = _bf(torch.randn(rows, ld, device=DEV, generator=g) * 3)
    logits = buf[:, :V]
    labels = torch.randint(1, V, (rows,), device=DEV, generator=g)
    labels[::7] = 0
    lr = logits.float().detach().clone().requires_grad_(True)
    mask = (labels != 0).float()
    lse = torch.logsumexp(lr, -1)
    loss = ((lse - lr.gather(-1, labels[:, None])[:, 0]) * mask).sum() / mask.sum() / workers
    loss.backward()
    correct = ((lr.argmax(-1) == labels) & (labels != 0)).sum()
    out3 = K.vocab_ce_fwd_bwd(buf, labels, V, workers, write_grad=True)
    assert abs(float(out3[0]) - float(loss)) < 2e-5 * max(1.0, abs(float(loss)))
    assert int(out3[1]) == int(correct) and int(out3[2]) == int(mask.sum())
    assert _rel(buf[:, :V], lr.grad) < 5e-3  # bf16 gradient
    assert torch.count_nonzero(buf[labels == 0][:, :V]) == 0


# ------------------------------------------------------------------------------------------ Adam
@pytest.mark.parametrize("iterations", [0, 1, 4000, 12345])
def test_adam_tf_form(iterations):
    from oracle import transformer_oracle as O
    g = torch.Generator(device=DEV).manual_seed(17)
    n = 1 << 16
    w = torch.randn(n, device=DEV, generator=g); gr = torch.randn(n, device=DEV, generator=g) * 1e-3
    gr[:100] = 1e-6
    m = torch.randn(n, device=DEV, generator=g) * 1e-3; v = torch.rand(n, device=DEV, generator=g) * 1e-6
    it = torch.tensor([iterations], device=DEV, dtype=torch.int64)
    lr = O.learning_rate(iterations, 128)
    we, me, ve = O.adam_update_numpy(w.cpu().numpy(), gr.cpu().numpy(), m.cpu().numpy(), v.cpu().numpy(), iterations, lr)
    w0 = w.clone()
    shadow = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    K.adam_step(w, gr, m, v, it, 128, w_bf16=shadow)
    if iterations == 0:
        assert torch.equal(w, w0)  # lr(0) == 0: weights bit-identical
    assert np.array_equal(m.cpu().numpy(), me) and np.array_equal(v.cpu().numpy(), ve)  # same IEEE op sequence: bit-exact
    upd, upd_e = (w - w0).cpu().numpy(), we - w0.cpu().numpy()
    np.testing.assert_allclose(w.cpu().numpy(), we, rtol=0, atol=2e-6 * float(np.abs(upd_e).max() + 1e-30) + 2.4e-7 * float(w0.abs().max()))
    assert torch.equal(shadow, w.to(torch.bfloat16))


# ------------------------------------------------------------------------------------------ fused GEMM epilogues (round 2)
@pytest.mark.parametrize("rows,d_in,d_out", [(4096, 512, 512), (4096, 2048, 512), (300, 128, 128), (77, 64, 64), (512, 1024, 1024),
                                              (130, 96, 200), (256, 256, 1536), (200, 64, 40)])
@pytest.mark.parametrize("with_resid", [True, False])
def test_dense_layernorm_fwd_fused(rows, d_in, d_out, with_resid):
    """z = resid + x W + b and y = LayerNorm(z) from ONE GEMM (cluster of up to 8 CTAs exchanging row statistics) against
    torch fp32: z bit-equal to the bf16 rounding of the fp32 expression up to accumulation order, y / mean / rstd with the
    tolerances of test_layernorm_fwd_bwd."""
    g = torch.Generator(device=DEV).manual_seed(rows + d_out)
    x = _bf(torch.randn(rows, d_in, device=DEV, generator=g)); w = _bf(torch.randn(d_in, d_out, device=DEV, generator=g) / math.sqrt(d_in))
    bias = torch.randn(d_out, device=DEV, generator=g); gamma = torch.randn(d_out, device=DEV, generator=g); beta = torch.randn(d_out, device=DEV, generator=g)
    resid = _bf(torch.randn(rows, d_out, device=DEV, generator=g)) if with_resid else None
    z, y, mean, rstd = K.dense_layernorm_fwd(x, w, bias, resid, gamma, beta, 1e-6)
    torch.cuda.synchronize()
    zr = x.float() @ w.float() + bias + (resid.float() if with_resid else 0.0)
    assert _rel(z, zr) < 4e-3 and float((z.float() - zr).abs().max()) < 6e-2          # one bf16 rounding
    zf = z.float()                                                                     # the statistics are those of the STORED z
    ref = torch.nn.functional.layer_norm(zf, (d_out,), gamma, beta, 1e-6)
    assert _rel(y, ref) < 4e-3 and float((y.float() - ref).abs().max()) < 6e-2
    np.testing.assert_allclose(mean.cpu(), zf.mean(-1).cpu(), atol=2e-5, rtol=1e-5)
    np.testing.assert_allclose(rstd.cpu(), (1.0 / torch.sqrt(zf.var(-1, unbiased=False) + 1e-6)).cpu(), rtol=2e-5)
    # and against the two-kernel path (GEMM, then the LayerNorm kernel) on the same inputs
    z2 = K.gemm(x, w, b_mn=True, bias=bias, resid=resid)
    y2, _, mean2, rstd2 = K.layernorm_residual_fwd(z2, None, gamma, beta, 1e-6)
    assert torch.equal(z2, z)
    assert _rel(y, y2) < 1e-3
    np.testing.assert_allclose(mean.cpu(), mean2.cpu(), atol=2e-6, rtol=1e-5)


@pytest.mark.parametrize("tokens,d_in,d_out", [(4096, 512, 1536), (4096, 512, 512), (4096, 512, 2048), (2560, 128, 384), (300, 64, 200),
                                               (4096, 512, 32000), (77, 40, 72)])
def test_dense_wgrad_with_fused_bias_gradient(tokens, d_in, d_out):
    """dW += x^T dy and dbias = sum_tokens dy from one weight-gradient GEMM (two extra warps sum the staged dy tiles; CTA
    pairs for the vocabulary-sized case) against torch fp32."""
    g = torch.Generator(device=DEV).manual_seed(tokens + d_out)
    ld = (d_out + 7) // 8 * 8
    x = _bf(torch.randn(tokens, d_in, device=DEV, generator=g))
    dy = _bf(torch.randn(tokens, ld, device=DEV, generator=g))[:, :d_out]
    dw0 = torch.randn(d_in, ld, device=DEV, generator=g)[:, :d_out]
    dw = dw0.clone()
    ldw = torch.zeros(d_in, ld, device=DEV); dwv = ldw[:, :d_out]; dwv.copy_(dw0)
    _, dbias = K.dense_wgrad_bias(x, dy, dwv)
    torch.cuda.synchronize()
    assert _rel(dwv, dw0 + x.float().t() @ dy.float()) < 2e-5
    assert _rel(dbias, dy.float().sum(0)) < 1e-5                                      # fp32 sums of exact bf16 values
