"""Synthetic token batches of the shapes BASELINE.json names: START=2, END=3, body ids
uniform in [4, V), PAD=0 right padding (PAD=0 is pinned by transformer_model.py:12,22,60)."""
from __future__ import annotations

import numpy as np

CONFIGS = {
    # name: hparams + batch geometry
    "tutorial": dict(layers_count=4, embedding_size=128, heads_count=8, feed_forward_hidden_size=512,
                     input_vocabulary_size=7765, target_vocabulary_size=7010, B=64, S=40, T=40),
    "base": dict(layers_count=6, embedding_size=512, heads_count=8, feed_forward_hidden_size=2048,
                 input_vocabulary_size=32000, target_vocabulary_size=32000, B=32, S=128, T=128),
    "big": dict(layers_count=6, embedding_size=1024, heads_count=16, feed_forward_hidden_size=4096,
                input_vocabulary_size=32000, target_vocabulary_size=32000, B=16, S=256, T=256),
    "long": dict(layers_count=6, embedding_size=512, heads_count=8, feed_forward_hidden_size=2048,
                 input_vocabulary_size=32000, target_vocabulary_size=32000, B=64, S=1024, T=1024),
}


def synthetic_batch(B, S, T_plus_1, vin, vtgt, seed, ragged=False):
    """Returns (input [B,S], target [B,T+1]) int64.  ragged=True draws lengths U{ceil(L/2)..L} and pads with 0."""
    rng = np.random.Generator(np.random.PCG64(seed))

    def make(n, L, V):
        out = np.zeros((n, L), dtype=np.int64)
        for b in range(n):
            ln = int(rng.integers((L + 1) // 2, L + 1)) if ragged else L
            ln = min(max(ln, 2), L)
            out[b, 0] = 2
            if L == 1:
                continue
            out[b, 1:ln - 1] = rng.integers(4, V, size=max(ln - 2, 0))
            out[b, ln - 1] = 3
        return out
    return make(B, S, vin), make(B, T_plus_1, vtgt)


def step_flops(L, d, dff, vtgt, Ns, Nt, S, T):
    """Algorithmic FLOPs of one train step: 3 x forward, dense attention, MAC = 2."""
    enc = L * (Ns * (8 * d * d + 4 * d * dff) + 4 * Ns * S * d)
    dec = L * (Nt * (8 * d * d + 4 * d * dff) + 4 * Nt * T * d + 4 * d * d * Nt + 4 * d * d * Ns + 4 * Nt * S * d)
    return 3 * (enc + dec + 2 * Nt * d * vtgt)
