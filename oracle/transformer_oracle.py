"""CPU fp32 restatement of the reference's data-parallel Transformer train_step.

TEST INFRASTRUCTURE ONLY.  Nothing under the product package may import this file; only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it,
and only as the checker / reported CPU baseline.

PARITY UNPINNED: the reference arithmetic lives in TensorFlow/Keras, which is un-vendored,
unpinned (reference setup.py:17-24, Dockerfile:1) and not installable here (no wheel, no
network); the reference ships no tests, golden vectors or fixtures for this path and its
checkpoint data shards are missing (.MISSING_LARGE_BLOBS).  What *is* pinned: the closed-form
known answers derivable from the reference source (tests/test_oracle_kat.py) and the
parameter names/shapes/offsets decoded from saved_weights/2/model_weights.index
(tests/golden/param_table.json).  The forward pass is also cross-checked against torch.nn's own encoder /
decoder stacks with the same weights (same test file): a guard against slips in the restatement, not a pin.

Every function cites the reference file:line it restates (paths are relative to
/root/reference/distributed_training_transformer/).  torch is used as a CPU fp32 array
library; autograd supplies an independent backward (the reference uses tf.GradientTape,
__main__.py:108-112).
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np
import torch

torch.set_grad_enabled(True)

NEG_MASK = -1e9  # transformer_model.py:101


# ----------------------------------------------------------------------------- closed forms
def positional_encoding(sentence_size: int, embedding_size: int) -> torch.Tensor:
    """transformer_model.py:29-53 — interleaved sin (even i) / cos (odd i), fp32, [1,P,d]."""
    pos = torch.arange(sentence_size, dtype=torch.float32).unsqueeze(-1)
    idx = torch.arange(embedding_size)
    expo = (2 * (idx // 2)).to(torch.float32) / torch.tensor(float(embedding_size), dtype=torch.float32)
    rates = 1.0 / torch.pow(torch.tensor(10000.0, dtype=torch.float32), expo)
    angles = pos * rates.unsqueeze(0)
    odd = (idx % 2).to(torch.float32).unsqueeze(0)
    even = 1.0 - odd
    return (odd * torch.cos(angles) + even * torch.sin(angles)).unsqueeze(0)


def create_padding_mask(sequence: torch.Tensor) -> torch.Tensor:
    """transformer_model.py:56-62 — 1.0 where token == 0, shape [B,1,1,S]."""
    return (sequence == 0).to(torch.float32)[:, None, None, :]


def create_look_ahead_mask(size: int) -> torch.Tensor:
    """transformer_model.py:65-70 — 1 - lower-triangular(ones) (strict upper triangle = 1)."""
    return 1.0 - torch.tril(torch.ones(size, size, dtype=torch.float32))


def create_masks(inp: torch.Tensor, target: torch.Tensor):
    """transformer_model.py:350-363 — (enc padding, max(target padding, look-ahead), dec padding=pad(input))."""
    enc = create_padding_mask(inp)
    dec = create_padding_mask(inp)
    la = create_look_ahead_mask(target.shape[1])
    la = torch.maximum(create_padding_mask(target), la)
    return enc, la, dec


def learning_rate(step, embedding_size: int, warmup_steps: int = 4000) -> np.float32:
    """__main__.py:56-67 — rsqrt(d) * min(step * warmup^-1.5, rsqrt(step)), all fp32.

    `step` is optimizer.iterations BEFORE the increment (legacy Keras OptimizerV2 passes
    float32(iterations) to the schedule), so the first step has lr == 0.
    """
    step = np.float32(step)
    with np.errstate(divide="ignore"):
        rising = step * np.float32(float(warmup_steps) ** -1.5)
        falling = np.float32(1.0) / np.sqrt(step, dtype=np.float32)
        return np.float32(np.float32(1.0) / np.sqrt(np.float32(embedding_size))) * np.minimum(rising, falling)


def adam_update_numpy(w, g, m, v, iterations: int, lr, beta1=0.9, beta2=0.98, eps=1e-9):
    """TF-form Adam (ResourceApplyAdam as driven by Keras OptimizerV2; __main__.py:72-73,113-114).

    t = iterations + 1; alpha = lr*sqrt(1-b2^t)/(1-b1^t); m += (g-m)(1-b1); v += (g*g-v)(1-b2);
    w -= alpha*m/(sqrt(v)+eps)  — epsilon is added to the un-bias-corrected sqrt(v).
    """
    f = np.float32
    t = f(iterations + 1)
    b1p = np.power(f(beta1), t, dtype=np.float32)
    b2p = np.power(f(beta2), t, dtype=np.float32)
    alpha = f(lr) * np.sqrt(f(1) - b2p, dtype=np.float32) / (f(1) - b1p)
    m = m + (g - m) * (f(1) - f(beta1))
    v = v + (g * g - v) * (f(1) - f(beta2))
    w = w - (m * alpha) / (np.sqrt(v, dtype=np.float32) + f(eps))
    return w.astype(np.float32), m.astype(np.float32), v.astype(np.float32)


# ----------------------------------------------------------------------------- parameters
def parameter_names(layers_count: int):
    """Keras object-graph names as in saved_weights/*/model_weights.index."""
    names = ["final_layer/kernel", "final_layer/bias",
             "encoder/embedding/embeddings", "decoder/embedding/embeddings"]
    for i in range(layers_count):
        for ln in ("layernorm1", "layernorm2"):
            names += [f"encoder/encoder_layers/{i}/{ln}/gamma", f"encoder/encoder_layers/{i}/{ln}/beta"]
    for i in range(layers_count):
        for ln in ("layernorm1", "layernorm2", "layernorm3"):
            names += [f"decoder/decoder_layers/{i}/{ln}/gamma", f"decoder/decoder_layers/{i}/{ln}/beta"]

    def mha(prefix):
        out = []
        for w in ("query_generator_weights", "key_generator_weights", "value_generator_weights", "dense"):
            out += [f"{prefix}/{w}/kernel", f"{prefix}/{w}/bias"]
        return out

    def ffn(prefix):
        return [f"{prefix}/ffn/layer_with_weights-0/kernel", f"{prefix}/ffn/layer_with_weights-0/bias",
                f"{prefix}/ffn/layer_with_weights-1/kernel", f"{prefix}/ffn/layer_with_weights-1/bias"]

    for i in range(layers_count):
        p = f"encoder/encoder_layers/{i}"
        names += mha(p + "/mha") + ffn(p)
    for i in range(layers_count):
        p = f"decoder/decoder_layers/{i}"
        names += mha(p + "/mha1") + mha(p + "/mha2") + ffn(p)
    return names


def parameter_shapes(layers_count, d, dff, vin, vtgt):
    shapes = OrderedDict()
    for n in parameter_names(layers_count):
        leaf = n.rsplit("/", 1)[1]
        if n == "final_layer/kernel": s = (d, vtgt)
        elif n == "final_layer/bias": s = (vtgt,)
        elif n == "encoder/embedding/embeddings": s = (vin, d)
        elif n == "decoder/embedding/embeddings": s = (vtgt, d)
        elif leaf in ("gamma", "beta"): s = (d,)
        elif "layer_with_weights-0" in n: s = (d, dff) if leaf == "kernel" else (dff,)
        elif "layer_with_weights-1" in n: s = (dff, d) if leaf == "kernel" else (d,)
        else: s = (d, d) if leaf == "kernel" else (d,)
        shapes[n] = s
    return shapes


def init_parameters(layers_count, d, dff, vin, vtgt, seed=0):
    """Keras default initialisers (transformer_model.py:119-122,172-174,259,290,333 pass none):
    Dense kernel Glorot-uniform, bias 0; Embedding U(-0.05,0.05); LayerNorm gamma 1, beta 0."""
    rng = np.random.Generator(np.random.PCG64(seed))
    out = OrderedDict()
    for n, s in parameter_shapes(layers_count, d, dff, vin, vtgt).items():
        leaf = n.rsplit("/", 1)[1]
        if leaf == "kernel":
            lim = math.sqrt(6.0 / (s[0] + s[1]))
            a = rng.uniform(-lim, lim, size=s)
        elif leaf == "embeddings": a = rng.uniform(-0.05, 0.05, size=s)
        elif leaf == "gamma": a = np.ones(s)
        else: a = np.zeros(s)
        out[n] = a.astype(np.float32)
    return out


# ----------------------------------------------------------------------------- model
def _round_bf16(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16).to(torch.float32)


class OracleTransformer:
    """transformer_model.py:318-363 restated over an explicit {name: tensor} parameter dict.

    emulate_bf16=True rounds GEMM operands to bf16 (fp32 accumulate) at the places the CUDA path
    does, to separate rounding noise from logic errors in tests; the default is the plain fp32
    restatement.
    """

    def __init__(self, layers_count, embedding_size, heads_count, feed_forward_hidden_size,
                 input_vocabulary_size, target_vocabulary_size, maximum_input_length,
                 maximum_target_length, droput_rate=0.0, emulate_bf16=False):
        assert embedding_size % heads_count == 0  # transformer_model.py:117
        # Dropout (transformer_model.py:190-191,223-225,268,299): TF's random stream cannot be reproduced, so with
        # rate > 0 the caller supplies the masks (values 1/(1-rate) or 0) per site in `dropout_masks`
        # {site id: [tokens, d] tensor}; site ids as in include/b200xf.h.  Without masks training=True is the identity.
        self.droput_rate = droput_rate
        self.dropout_masks = {}
        self.L, self.d, self.H, self.dff = layers_count, embedding_size, heads_count, feed_forward_hidden_size
        self.vin, self.vtgt = input_vocabulary_size, target_vocabulary_size
        self.pe_in = positional_encoding(maximum_input_length, embedding_size)
        self.pe_tgt = positional_encoding(maximum_target_length, embedding_size)
        self.emulate_bf16 = emulate_bf16
        self.params: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        self.m: "OrderedDict[str, np.ndarray]" = OrderedDict()
        self.v: "OrderedDict[str, np.ndarray]" = OrderedDict()
        self.iterations = 0
        self.warmup_steps = 4000

    def load(self, arrays):
        for n in parameter_names(self.L):
            self.params[n] = torch.tensor(np.asarray(arrays[n], dtype=np.float32), requires_grad=True)
            self.m[n] = np.zeros_like(arrays[n], dtype=np.float32)
            self.v[n] = np.zeros_like(arrays[n], dtype=np.float32)
        return self

    # -- layers ---------------------------------------------------------------------------
    def _r(self, x):
        return _round_bf16(x) if self.emulate_bf16 else x

    def _dense(self, x, prefix):
        """tf.keras.layers.Dense: y = x @ kernel[in,out] + bias."""
        return self._r(x) @ self._r(self.params[prefix + "/kernel"]) + self.params[prefix + "/bias"]

    @staticmethod
    def _layernorm(x, gamma, beta, eps=1e-6):
        """LayerNormalization(epsilon=1e-6) over the last axis, biased variance (transformer_model.py:187-188)."""
        mean = x.mean(-1, keepdim=True)
        var = ((x - mean) ** 2).mean(-1, keepdim=True)
        return (x - mean) * torch.rsqrt(var + eps) * gamma + beta

    def _sdpa(self, q, k, v, mask):
        """transformer_model.py:73-109."""
        logits = self._r(q) @ self._r(k).transpose(-1, -2)
        logits = logits / math.sqrt(k.shape[-1])
        if mask is not None:
            logits = logits + mask * NEG_MASK
        w = torch.softmax(logits, dim=-1)
        return self._r(w) @ self._r(v), w

    def _mha(self, values, keys, query, mask, prefix):
        """transformer_model.py:133-166 — NOTE the (values, keys, query, mask) argument order."""
        B = query.shape[0]
        q = self._dense(query, prefix + "/query_generator_weights")
        k = self._dense(keys, prefix + "/key_generator_weights")
        v = self._dense(values, prefix + "/value_generator_weights")
        split = lambda x: x.reshape(B, -1, self.H, self.d // self.H).permute(0, 2, 1, 3)
        o, w = self._sdpa(split(q), split(k), split(v), mask)
        o = o.permute(0, 2, 1, 3).reshape(B, -1, self.d)
        return self._dense(o, prefix + "/dense"), w

    def _ffn(self, x, prefix):
        """transformer_model.py:169-175."""
        h = torch.relu(self._dense(x, prefix + "/ffn/layer_with_weights-0"))
        return self._dense(h, prefix + "/ffn/layer_with_weights-1")

    def _ln(self, x, prefix):
        return self._layernorm(x, self.params[prefix + "/gamma"], self.params[prefix + "/beta"])

    def _drop(self, x, site, training):
        """tf.keras.layers.Dropout with an externally supplied mask (see __init__)."""
        if not training or site not in self.dropout_masks:
            return x
        return x * self.dropout_masks[site].reshape(x.shape)

    def encoder(self, tokens, mask, training=False):
        """transformer_model.py:270-279."""
        S = tokens.shape[1]
        x = self.params["encoder/embedding/embeddings"][tokens]
        x = x * math.sqrt(float(self.d))
        x = x + self.pe_in[:, :S, :]
        x = self._drop(x, 0, training)                                                    # :276
        for i in range(self.L):
            p = f"encoder/encoder_layers/{i}"
            a, _ = self._mha(x, x, x, mask, p + "/mha")                                   # :195
            o1 = self._ln(x + self._drop(a, 2 + 2 * i, training), p + "/layernorm1")      # :196-198
            x = self._ln(o1 + self._drop(self._ffn(o1, p), 3 + 2 * i, training), p + "/layernorm2")  # :201-204
        return x

    def decoder(self, tokens, enc_out, la_mask, pad_mask, training=False):
        """transformer_model.py:301-315."""
        T = tokens.shape[1]
        weights = {}
        x = self.params["decoder/embedding/embeddings"][tokens]
        x = x * math.sqrt(float(self.d))
        x = x + self.pe_tgt[:, :T, :]
        x = self._drop(x, 1, training)                                                # :308
        base = 2 + 2 * self.L
        for i in range(self.L):
            p = f"decoder/decoder_layers/{i}"
            a1, w1 = self._mha(x, x, x, la_mask, p + "/mha1")                       # :230
            o1 = self._ln(self._drop(a1, base + 3 * i, training) + x, p + "/layernorm1")              # :232-233
            a2, w2 = self._mha(enc_out, enc_out, o1, pad_mask, p + "/mha2")          # :236-237
            o2 = self._ln(self._drop(a2, base + 3 * i + 1, training) + o1, p + "/layernorm2")         # :238-240
            x = self._ln(self._drop(self._ffn(o2, p), base + 3 * i + 2, training) + o2, p + "/layernorm3")  # :243-246
            weights[f"decoder_layer{i+1}_block1"] = w1
            weights[f"decoder_layer{i+1}_block2"] = w2
        return x, weights

    def __call__(self, inputs, training=False):
        """transformer_model.py:335-348 — returns (logits [B,T,V], attention-weights dict)."""
        inp, tgt = inputs
        enc_mask, la_mask, dec_mask = create_masks(inp, tgt)
        enc_out = self.encoder(inp, enc_mask, training)
        dec_out, w = self.decoder(tgt, enc_out, la_mask, dec_mask, training)
        return self._dense(dec_out, "final_layer"), w


# ----------------------------------------------------------------------------- loss / step
def local_loss_function(actual, predicted, workers_count):
    """transformer_model.py:11-17 — sum(masked CE)/sum(mask)/workers_count."""
    mask = (actual != 0)
    lse = torch.logsumexp(predicted, dim=-1)
    picked = predicted.gather(-1, actual.unsqueeze(-1)).squeeze(-1)
    loss_ = (lse - picked) * mask.to(predicted.dtype)
    return loss_.sum() / mask.to(predicted.dtype).sum() / workers_count


def accuracy_function(actual, predicted):
    """transformer_model.py:20-26 — masked mean of argmax == label (first max wins)."""
    mask = (actual != 0)
    pred = _argmax_first(predicted)
    acc = (pred == actual) & mask
    return acc.to(torch.float32).sum() / mask.to(torch.float32).sum()


def _argmax_first(x):
    mx = x.max(dim=-1, keepdim=True).values
    idx = torch.arange(x.shape[-1]).expand_as(x)
    big = torch.full_like(idx, x.shape[-1])
    return torch.where(x == mx, idx, big).min(dim=-1).values


def accuracy_counts(actual, predicted):
    mask = (actual != 0)
    pred = _argmax_first(predicted)
    return int(((pred == actual) & mask).sum()), int(mask.sum())


def train_step(model: OracleTransformer, inp, target, workers_count=1, allreduce=None):
    """__main__.py:105-117.  `allreduce(grads: dict)` (optional) emulates the implicit SUM over
    replicas inside apply_gradients (cluster/cluster.py:66).  Returns (loss, logits, grads)."""
    inp = torch.as_tensor(inp, dtype=torch.int64); target = torch.as_tensor(target, dtype=torch.int64)
    target_input, target_real = target[:, :-1], target[:, 1:]
    for p in model.params.values():
        p.grad = None
    logits, _ = model([inp, target_input], training=True)
    loss = local_loss_function(target_real, logits, workers_count)
    loss.backward()
    grads = OrderedDict((n, p.grad.detach().numpy().copy()) for n, p in model.params.items())
    if allreduce is not None:
        grads = allreduce(grads)
    lr = learning_rate(model.iterations, model.d, model.warmup_steps)
    with torch.no_grad():
        for n, p in model.params.items():
            w, m, v = adam_update_numpy(p.detach().numpy(), grads[n], model.m[n], model.v[n], model.iterations, lr)
            p.copy_(torch.from_numpy(w)); model.m[n] = m; model.v[n] = v
    model.iterations += 1
    return float(loss.detach()), logits.detach(), grads


# ----------------------------------------------------------------------------- synthetic data
def synthetic_batch(B, S, T_plus_1, vin, vtgt, seed, ragged=True):
    """START=2, END=3, body ids U[4,V), PAD=0 right padding; PCG64(seed)."""
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
