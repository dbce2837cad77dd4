"""torch-tensor wrappers over the per-kernel C entry points (used by tests, ncu runs and bench.py).

torch is only the allocator / stream provider here: every function forwards raw device pointers to
libb200xf.so and launches on torch's current CUDA stream.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.B200XFError("b200xf kernels need CUDA tensors: there is no CPU fallback")


def gemm(a, b, *, a_mn=False, b_mn=False, bias=None, relu=False, resid=None, resid_op=1, out=None,
         out_fp32=False, accumulate=False, alpha=1.0, force_bn=0, split_k=0):
    """D[M,N] = alpha * sum_k A[m,k] B[n,k] (+bias)(relu)(+resid | relu-mask(resid)).

    a: [M,K] (a_mn False) or [K,M] (a_mn True); b: [N,K] (b_mn False) or [K,N] (b_mn True); bf16, row-major.
    """
    _need_cuda(a, b)
    lib = _lib.load()
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N = b.shape[1] if b_mn else b.shape[0]
    assert (b.shape[0] if b_mn else b.shape[1]) == K
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=torch.float32 if out_fp32 else torch.bfloat16)
        if accumulate:
            out.zero_()
    check(lib.b200xf_gemm_bf16(ptr(a), a.stride(0), int(a_mn), ptr(b), b.stride(0), int(b_mn), ptr(out), out.stride(0),
                               int(out.dtype == torch.float32), int(accumulate), M, N, K, ptr(bias), int(relu),
                               ptr(resid), resid.stride(0) if resid is not None else 0,
                               resid_op if resid is not None else 0, float(alpha), force_bn, split_k, stream_ptr()))
    return out


def dense_layernorm_fwd(x, w, bias, resid, gamma, beta, eps=1e-6):
    """z = resid + x @ w + bias (bf16), y = LayerNorm(z): one GEMM with the normalisation in its epilogue.
    x [rows, d_in] bf16, w [d_in, d_out] bf16; returns (z, y, mean, rstd)."""
    _need_cuda(x, w, gamma, beta)
    rows, d_in = x.shape
    d_out = w.shape[1]
    z = torch.empty((rows, d_out), device=x.device, dtype=torch.bfloat16); y = torch.empty_like(z)
    mean = torch.empty(rows, device=x.device, dtype=torch.float32); rstd = torch.empty_like(mean)
    check(_lib.load().b200xf_dense_layernorm_fwd(ptr(x), x.stride(0), ptr(w), w.stride(0), ptr(bias), ptr(resid),
                                                 resid.stride(0) if resid is not None else 0, ptr(z), z.stride(0), ptr(gamma), ptr(beta),
                                                 ptr(y), y.stride(0), ptr(mean), ptr(rstd), rows, d_out, d_in, float(eps), stream_ptr()))
    return z, y, mean, rstd


def dense_wgrad_bias(x, dy, dw=None):
    """dw += x^T dy (fp32), dbias = column sums of dy, from one pass over dy.  x [tokens, d_in], dy [tokens, d_out] bf16."""
    _need_cuda(x, dy)
    tokens, d_in = x.shape
    d_out = dy.shape[1]
    if dw is None:
        dw = torch.zeros((d_in, d_out), device=x.device, dtype=torch.float32)
    dbias = torch.empty(d_out, device=x.device, dtype=torch.float32)
    scratch = torch.empty(148 * d_out, device=x.device, dtype=torch.float32)
    check(_lib.load().b200xf_dense_wgrad_bias(ptr(x), x.stride(0), ptr(dy), dy.stride(0), ptr(dw), dw.stride(0), ptr(dbias), ptr(scratch),
                                              scratch.numel(), tokens, d_in, d_out, stream_ptr()))
    return dw, dbias


def attention_fwd(q, k, v, key_tokens, causal, H):
    """q [B,Sq,H*dk], k/v [B,Sk,H*dk] bf16 (may be column slices of wider buffers); returns (o, lse)."""
    _need_cuda(q, k, v, key_tokens)
    lib = _lib.load()
    B, Sq, D = q.shape
    Sk = k.shape[1]
    dk = D // H
    o = torch.empty((B, Sq, D), device=q.device, dtype=torch.bfloat16)
    lse = torch.empty((B, H, Sq, 2), device=q.device, dtype=torch.float32)
    check(lib.b200xf_attention_fwd(ptr(q), q.stride(1), ptr(k), k.stride(1), ptr(v), v.stride(1), ptr(o), o.stride(1), ptr(lse),
                                   ptr(key_tokens), int(causal), B, H, Sq, Sk, dk, stream_ptr()))
    return o, lse


def attention_bwd(q, k, v, o, lse, d_o, key_tokens, causal, H, dq_scratch=True):
    _need_cuda(q, k, v, o, d_o)
    lib = _lib.load()
    B, Sq, D = q.shape
    Sk = k.shape[1]
    dk = D // H
    dq = torch.empty((B, Sq, D), device=q.device, dtype=torch.bfloat16)
    dkk = torch.empty((B, Sk, D), device=q.device, dtype=torch.bfloat16)
    dv = torch.empty((B, Sk, D), device=q.device, dtype=torch.bfloat16)
    delta = torch.empty((B, H, Sq), device=q.device, dtype=torch.float32)
    dq_acc = torch.empty((B * Sq, D), device=q.device, dtype=torch.float32) if dq_scratch else None
    check(lib.b200xf_attention_bwd(ptr(q), q.stride(1), ptr(k), k.stride(1), ptr(v), v.stride(1), ptr(o), o.stride(1), ptr(lse),
                                   ptr(key_tokens), int(causal), B, H, Sq, Sk, dk, ptr(d_o), d_o.stride(1), ptr(dq), dq.stride(1),
                                   ptr(dkk), dkk.stride(1), ptr(dv), dv.stride(1), ptr(delta), ptr(dq_acc), stream_ptr()))
    return dq, dkk, dv


def layernorm_residual_fwd(x, resid, gamma, beta, eps=1e-6, want_z=False):
    _need_cuda(x, gamma, beta)
    lib = _lib.load()
    rows, d = x.shape
    y = torch.empty_like(x)
    z = torch.empty_like(x) if (want_z and resid is not None) else None
    mean = torch.empty(rows, device=x.device, dtype=torch.float32)
    rstd = torch.empty(rows, device=x.device, dtype=torch.float32)
    check(lib.b200xf_layernorm_residual_fwd(ptr(x), ptr(resid), ptr(gamma), ptr(beta), ptr(y), ptr(z), ptr(mean), ptr(rstd),
                                            rows, d, float(eps), stream_ptr()))
    return y, z, mean, rstd


def layernorm_bwd_scratch(rows, d, device):
    nbytes = _lib.load().b200xf_layernorm_bwd_scratch_bytes(rows, d)
    return torch.zeros((nbytes + 3) // 4, device=device, dtype=torch.int32)


def layernorm_residual_bwd(dy, z, mean, rstd, gamma, scratch=None, want_dcol=True):
    _need_cuda(dy, z)
    lib = _lib.load()
    rows, d = dy.shape
    if scratch is None:
        scratch = layernorm_bwd_scratch(rows, d, dy.device)
    dz = torch.empty_like(dy)
    dgamma = torch.empty(d, device=dy.device, dtype=torch.float32)
    dbeta = torch.empty(d, device=dy.device, dtype=torch.float32)
    dcol = torch.empty(d, device=dy.device, dtype=torch.float32) if want_dcol else None
    check(lib.b200xf_layernorm_residual_bwd(ptr(dy), ptr(z), ptr(mean), ptr(rstd), ptr(gamma), ptr(dz), ptr(dgamma), ptr(dbeta),
                                            ptr(dcol), rows, d, ptr(scratch), stream_ptr()))
    return dz, dgamma, dbeta, dcol


def positional_encoding(positions, d):
    out = torch.empty((positions, d), dtype=torch.float32)
    check(_lib.load().b200xf_positional_encoding_host(C.c_void_p(out.data_ptr()), positions, d))
    return out


def embed_pe_fwd(tokens, table, pe):
    _need_cuda(tokens, table, pe)
    B, S = tokens.shape
    V, d = table.shape
    out = torch.empty((B, S, d), device=table.device, dtype=torch.bfloat16)
    check(_lib.load().b200xf_embed_pe_fwd(ptr(tokens), ptr(table), V, ptr(pe), ptr(out), B, S, d, stream_ptr()))
    return out


def embed_pe_bwd(tokens, dx, vocab):
    _need_cuda(tokens, dx)
    lib = _lib.load()
    n = tokens.numel()
    d = dx.shape[-1]
    dtable = torch.zeros((vocab, d), device=dx.device, dtype=torch.float32)
    nbytes = lib.b200xf_embed_bwd_scratch_bytes(n)
    scratch = torch.empty(nbytes, device=dx.device, dtype=torch.uint8)
    check(lib.b200xf_embed_pe_bwd(ptr(tokens), ptr(dx), ptr(dtable), vocab, n, d, ptr(scratch), nbytes, stream_ptr()))
    return dtable


def vocab_ce_fwd_bwd(logits, labels, V, workers_count=1, write_grad=True):
    """logits bf16 [rows, ld>=V] (overwritten with dlogits when write_grad); returns tensor [loss, acc_num, acc_den]."""
    _need_cuda(logits, labels)
    rows = logits.shape[0]
    out3 = torch.empty(3, device=logits.device, dtype=torch.float32)
    scratch = torch.empty(rows * 2 + 4, device=logits.device, dtype=torch.float32)
    check(_lib.load().b200xf_vocab_ce_fwd_bwd(ptr(logits), logits.stride(0), ptr(labels), rows, V, workers_count, int(write_grad),
                                              ptr(out3), ptr(scratch), stream_ptr()))
    return out3


def adam_step(w, g, m, v, iterations_dev, embedding_size, warmup_steps=4000, beta_1=0.9, beta_2=0.98, epsilon=1e-9,
              w_bf16=None, zero_grad=False):
    _need_cuda(w, g, m, v, iterations_dev)
    check(_lib.load().b200xf_adam_step(ptr(w), ptr(g), ptr(m), ptr(v), ptr(w_bf16), w.numel(), ptr(iterations_dev), embedding_size,
                                       warmup_steps, beta_1, beta_2, epsilon, int(zero_grad), stream_ptr()))
