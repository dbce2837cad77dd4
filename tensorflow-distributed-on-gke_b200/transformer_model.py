"""Host-side mirror of the reference's `transformer_model.py` public surface over libb200xf.so.

Same names, argument order and meaning as /root/reference/distributed_training_transformer/
transformer_model.py (cited per symbol); the arithmetic runs in the CUDA library.  The tutorial
names that BASELINE.json uses (num_layers, d_model, ...) are accepted as keyword aliases.
"""
from __future__ import annotations

import ctypes as C
import math
from collections import OrderedDict

import numpy as np
import torch

from . import _lib
from ._lib import check, ptr

_ALIASES = {  # TF-tutorial keyword -> reference keyword
    "num_layers": "layers_count", "d_model": "embedding_size", "num_heads": "heads_count",
    "dff": "feed_forward_hidden_size", "input_vocab_size": "input_vocabulary_size",
    "target_vocab_size": "target_vocabulary_size", "pe_input": "maximum_input_length",
    "pe_target": "maximum_target_length", "rate": "droput_rate", "dropout_rate": "droput_rate",
}


def _cuda_tokens(x, device):
    t = torch.as_tensor(np.asarray(x) if not isinstance(x, torch.Tensor) else x)
    return t.to(device=device, dtype=torch.int64).contiguous()


# ------------------------------------------------------------------ masks / PE (host logic, exact)
def create_padding_mask(sequence):
    """transformer_model.py:56-62 — 1.0 where token == 0, shape [B,1,1,S]."""
    s = torch.as_tensor(sequence)
    return (s == 0).to(torch.float32)[:, None, None, :]


def create_look_ahead_mask(size):
    """transformer_model.py:65-70 — strict upper triangle = 1."""
    return 1.0 - torch.tril(torch.ones((int(size), int(size)), dtype=torch.float32))


def positional_encoding(sentence_size, embedding_size):
    """transformer_model.py:29-53 — [1, P, d] fp32, sin on even / cos on odd columns (computed by the library)."""
    from .kernels import positional_encoding as pe
    return pe(int(sentence_size), int(embedding_size)).unsqueeze(0)


# ------------------------------------------------------------------ variables
class Variable:
    """A trainable variable: a strided fp32 view into the library's flat parameter buffer."""

    def __init__(self, model, name, index, rows, cols, ld, offset, ndim):
        self._model, self.name, self.index = model, name, index
        self.rows, self.cols, self.ld, self.offset, self.ndim = rows, cols, ld, offset, ndim
        self.shape = (rows, cols) if ndim == 2 else (cols,)

    def _view(self, flat):
        v = flat.as_strided((self.rows, self.cols), (self.ld, 1), self.offset)
        return v if self.ndim == 2 else v[0]

    @property
    def value(self):
        return self._view(self._model._flat("param"))

    @property
    def gradient(self):
        return self._view(self._model._flat("grad"))

    def numpy(self):
        return self.value.detach().cpu().numpy().copy()

    def assign(self, array):
        self.value.copy_(torch.as_tensor(np.asarray(array, dtype=np.float32)).to(self.value.device).reshape(self.shape))
        self._model._dirty = True


class _DeviceArray:
    """Minimal __cuda_array_interface__ holder so torch can wrap a borrowed device pointer without copying."""

    def __init__(self, address, count, owner):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": "<f4", "data": (address, False), "version": 3}
        self._owner = owner


class AttentionWeights(dict):
    """The attention_weights dict of Transformer.call (transformer_model.py:312-313): keys
    `decoder_layer{i}_block1` ([B,H,T,T]) and `decoder_layer{i}_block2` ([B,H,T,S]), i = 1..layers_count.  The maps are
    computed on first access from the activations of the forward pass that returned this object, so read them before
    the model runs another batch."""

    def __init__(self, model, inp, tgt):
        super().__init__()
        self._model, self._inp, self._tgt = model, inp, tgt   # keeps the token buffers alive
        self._serial = model._forward_serial
        self._names = [f"decoder_layer{i + 1}_block{b}" for i in range(model.hparams["layers_count"]) for b in (1, 2)]

    def keys(self):
        return list(self._names)

    def __iter__(self):
        return iter(self._names)

    def __len__(self):
        return len(self._names)

    def __contains__(self, k):
        return k in self._names

    def items(self):
        return [(k, self[k]) for k in self._names]

    def values(self):
        return [self[k] for k in self._names]

    def __getitem__(self, key):
        if dict.__contains__(self, key):
            return dict.__getitem__(self, key)
        if key not in self._names:
            raise KeyError(key)
        m = self._model
        if m._forward_serial != self._serial:
            raise _lib.B200XFError("attention weights must be read before the model runs another batch")
        layer = int(key.split("_")[1][len("layer"):]) - 1
        block = int(key[-1])
        B, S = self._inp.shape; T = self._tgt.shape[1]
        out = torch.empty((B, m.hparams["heads_count"], T, T if block == 1 else S), device=m.device, dtype=torch.float32)
        check(_lib.load().b200xf_attention_weights(m._ctx, layer, block, ptr(out)))
        dict.__setitem__(self, key, out)
        return out


# ------------------------------------------------------------------ the model
class Transformer:
    """transformer_model.py:318-363 — Transformer(layers_count, embedding_size, heads_count,
    feed_forward_hidden_size, input_vocabulary_size, target_vocabulary_size, maximum_input_length,
    maximum_target_length, droput_rate=0.1).

    Extra keyword-only arguments (no counterpart in the reference): `max_tokens` (initial workspace
    capacity in tokens per side; grows on demand), `device`, `seed` (Keras-default initialisers),
    optimizer constants used by the fused Adam (`adam_beta_1`, `adam_beta_2`, `adam_epsilon`, `warmup_steps`).
    """

    def __init__(self, layers_count=None, embedding_size=None, heads_count=None, feed_forward_hidden_size=None,
                 input_vocabulary_size=None, target_vocabulary_size=None, maximum_input_length=None,
                 maximum_target_length=None, droput_rate=0.1, *, max_tokens=4096, device=None, seed=0,
                 adam_beta_1=0.9, adam_beta_2=0.98, adam_epsilon=1e-9, warmup_steps=4000, **aliases):
        args = dict(layers_count=layers_count, embedding_size=embedding_size, heads_count=heads_count,
                    feed_forward_hidden_size=feed_forward_hidden_size, input_vocabulary_size=input_vocabulary_size,
                    target_vocabulary_size=target_vocabulary_size, maximum_input_length=maximum_input_length,
                    maximum_target_length=maximum_target_length, droput_rate=droput_rate)
        for k, v in aliases.items():
            if k not in _ALIASES:
                raise TypeError(f"Transformer() got an unexpected keyword argument {k!r}")
            args[_ALIASES[k]] = v
        missing = [k for k, v in args.items() if v is None]
        if missing:
            raise TypeError(f"Transformer() missing arguments: {missing}")
        if int(args["embedding_size"]) % int(args["heads_count"]) != 0:
            raise AssertionError("embedding_size % heads_count != 0")  # transformer_model.py:117
        self.hparams = {k: (float(v) if k == "droput_rate" else int(v)) for k, v in args.items()}
        self._opt = dict(adam_beta_1=float(adam_beta_1), adam_beta_2=float(adam_beta_2), adam_epsilon=float(adam_epsilon),
                         warmup_steps=int(warmup_steps))
        self._max_tokens = int(max_tokens)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else None
        self.device = torch.device(device) if device is not None else None
        self._seed = seed
        self._ctx = None
        self._flats = {}
        self._dirty = False
        self.variables = []
        self._comm = None
        self._symmetric = None   # symmetric-memory allocations behind weights / gradients once the NVLS update is attached
        self._forward_serial = 0

    # ---- context management ---------------------------------------------------------------
    def _create_ctx(self, max_tokens):
        if self.device is None or self.device.type != "cuda":
            raise _lib.B200XFError("b200xf needs a CUDA device (sm_100a): there is no CPU fallback")
        lib = _lib.load()
        hp = _lib.HParams(**self.hparams, **self._opt, max_tokens=int(max_tokens))
        ctx = C.c_void_p()
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream().cuda_stream
            check(lib.b200xf_ctx_create(C.byref(hp), self.device.index or 0, C.c_void_p(stream), C.byref(ctx)))
        return ctx

    def _bind(self, ctx):
        lib = _lib.load()
        self._ctx = ctx
        total = lib.b200xf_param_total(ctx)
        self._flats = {}
        for kind, fn in (("param", lib.b200xf_param_ptr), ("grad", lib.b200xf_grad_ptr), ("adam_m", lib.b200xf_adam_m_ptr),
                         ("adam_v", lib.b200xf_adam_v_ptr)):
            self._flats[kind] = torch.as_tensor(_DeviceArray(fn(ctx), total, self), device=self.device)
        self.variables = []
        name = C.create_string_buffer(256)
        rows, cols, ld, off, nd = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64(), C.c_int32()
        for i in range(lib.b200xf_param_count(ctx)):
            check(lib.b200xf_param_info(ctx, i, name, 256, C.byref(rows), C.byref(cols), C.byref(ld), C.byref(off), C.byref(nd)))
            self.variables.append(Variable(self, name.value.decode(), i, rows.value, cols.value, ld.value, off.value, nd.value))

    def _flat(self, kind):
        self._ensure_built()
        return self._flats[kind]

    def _ensure_built(self, need_tokens=0):
        if self._ctx is None:
            self._max_tokens = max(self._max_tokens, int(need_tokens))
            self._bind(self._create_ctx(self._max_tokens))
            self._init_weights()
        elif need_tokens > self._max_tokens:
            self._grow(int(need_tokens))

    def _grow(self, need_tokens):
        """Re-create the workspace with a larger token capacity, carrying weights and optimizer state over."""
        if self._comm is not None:
            raise _lib.B200XFError("workspace growth after comm_init is not supported: construct with max_tokens large enough")
        lib = _lib.load()
        old_ctx, old = self._ctx, {k: v.clone() for k, v in self._flats.items()}
        it = self.iterations
        new_cap = max(need_tokens, 2 * self._max_tokens)
        self._flats = {}
        check(lib.b200xf_ctx_destroy(old_ctx))
        self._bind(self._create_ctx(new_cap))
        self._max_tokens = new_cap
        for k, v in old.items():
            self._flats[k].copy_(v)
        check(lib.b200xf_params_changed(self._ctx))
        self.iterations = it

    def _init_weights(self):
        """Keras defaults (the reference passes no initialisers: transformer_model.py:119-122,172-174,259,290,333):
        Dense kernel Glorot-uniform, bias 0, Embedding U(-0.05, 0.05), LayerNorm gamma 1 / beta 0."""
        rng = np.random.Generator(np.random.PCG64(self._seed))
        self._flats["param"].zero_()
        for v in self.variables:
            leaf = v.name.rsplit("/", 1)[1]
            if leaf == "kernel":
                lim = math.sqrt(6.0 / (v.shape[0] + v.shape[1]))
                a = rng.uniform(-lim, lim, size=v.shape)
            elif leaf == "embeddings":
                a = rng.uniform(-0.05, 0.05, size=v.shape)
            elif leaf == "gamma":
                a = np.ones(v.shape)
            else:
                a = np.zeros(v.shape)
            v.assign(a.astype(np.float32))
        self._sync_weights()

    def _sync_weights(self):
        if self._dirty and self._ctx is not None:
            check(_lib.load().b200xf_params_changed(self._ctx))
            self._dirty = False

    def close(self):
        """Release the device context (weights, optimizer state, workspace)."""
        if self._ctx is not None:
            self._flats = {}
            _lib.load().b200xf_ctx_destroy(self._ctx)
            self._ctx = None
            self._symmetric = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- Keras-like surface ----------------------------------------------------------------------
    @property
    def trainable_variables(self):
        self._ensure_built()
        return list(self.variables)

    @property
    def iterations(self):
        self._ensure_built()
        v = C.c_int64()
        check(_lib.load().b200xf_get_iterations(self._ctx, C.byref(v)))
        return v.value

    @iterations.setter
    def iterations(self, value):
        self._ensure_built()
        check(_lib.load().b200xf_set_iterations(self._ctx, int(value)))

    def get_weights_dict(self):
        self._ensure_built()
        return OrderedDict((v.name, v.numpy()) for v in self.variables)

    def set_weights_dict(self, arrays):
        self._ensure_built()
        for v in self.variables:
            v.assign(arrays[v.name])
        self._sync_weights()

    def gather_optimizer_state(self):
        """Collective (every replica calls it): after sharded multi-replica steps each replica holds current Adam slots
        only for its shard of every bucket; this all-gathers them so that every replica holds all of m and v."""
        self._ensure_built()
        check(_lib.load().b200xf_gather_optimizer_state(self._ctx))

    def get_slot_dict(self, kind):
        """Adam slots (`adam_m` / `adam_v`) or `grad`, as {name: ndarray}."""
        flat = self._flat(kind)
        if kind in ("adam_m", "adam_v") and not _lib.load().b200xf_optimizer_state_complete(self._ctx):
            raise _lib.B200XFError(f"{kind} is sharded across replicas after a multi-replica step: call "
                                   "gather_optimizer_state() on every replica first")
        return OrderedDict((v.name, v._view(flat).detach().cpu().numpy().copy()) for v in self.variables)

    def create_masks(self, input, target):
        """transformer_model.py:350-363 — (encoder padding, max(target padding, look-ahead), decoder padding = pad(input))."""
        enc = create_padding_mask(input)
        dec = create_padding_mask(input)
        la = create_look_ahead_mask(torch.as_tensor(target).shape[1])
        la = torch.maximum(create_padding_mask(target), la)
        return enc, la, dec

    def __call__(self, inputs, training=False):
        """transformer_model.py:335-348 — returns (logits fp32 [B,T,V] on the device, attention_weights dict).
        The masks of create_masks are applied inside the attention kernels from the token ids."""
        inp, tgt = inputs
        inp = _cuda_tokens(inp, self.device); tgt = _cuda_tokens(tgt, self.device)
        B, S = inp.shape
        T = tgt.shape[1]
        self._ensure_built(max(B * S, B * T))
        self._sync_weights()
        self._forward_serial += 1
        logits = torch.empty((B, T, self.hparams["target_vocabulary_size"]), device=self.device, dtype=torch.float32)
        check(_lib.load().b200xf_forward(self._ctx, ptr(inp), ptr(tgt), B, S, T, int(bool(training)), ptr(logits)))
        return logits, AttentionWeights(self, inp, tgt)

    call = __call__

    # ---- fused steps ---------------------------------------------------------------------------------
    def _step(self, fn, inp, target, *extra):
        inp = _cuda_tokens(inp, self.device); target = _cuda_tokens(target, self.device)
        B, S = inp.shape
        T1 = target.shape[1]
        self._ensure_built(max(B * S, B * T1))
        self._sync_weights()
        self._forward_serial += 1
        out = torch.empty(6, device=self.device, dtype=torch.float32)
        check(fn(self._ctx, ptr(inp), ptr(target), B, S, T1, *extra, ptr(out)))
        return out

    def train_step_device(self, inp, target, workers_count=1):
        """__main__.py:105-117 in one call; returns a device tensor [loss, acc_num, acc_den, sum_loss, sum_num, sum_den]."""
        return self._step(_lib.load().b200xf_train_step_device, inp, target, int(workers_count))

    def train_step_host(self, inp, target, workers_count=1):
        """Same step through the HOST-buffer entry point: `inp`/`target` are CPU int64 tensors (pinned for speed);
        the copy in, the step and the copy of the scalars back all happen inside the call."""
        inp = torch.as_tensor(inp, dtype=torch.int64).contiguous(); target = torch.as_tensor(target, dtype=torch.int64).contiguous()
        assert not inp.is_cuda and not target.is_cuda
        B, S = inp.shape
        T1 = target.shape[1]
        self._ensure_built(max(B * S, B * T1))
        self._sync_weights()
        self._forward_serial += 1
        out = (C.c_float * 6)()
        check(_lib.load().b200xf_train_step_host(self._ctx, C.c_void_p(inp.data_ptr()), C.c_void_p(target.data_ptr()), B, S, T1,
                                                 int(workers_count), out))
        return list(out)

    def eval_step_device(self, inp, target):
        """test_step, __main__.py:119-125."""
        return self._step(_lib.load().b200xf_eval_step_device, inp, target)

    def forward_backward_device(self, inp, target, workers_count=1):
        return self._step(_lib.load().b200xf_forward_backward_device, inp, target, int(workers_count))

    def set_dropout_seed(self, seed):
        self._ensure_built()
        check(_lib.load().b200xf_set_dropout_seed(self._ctx, int(seed)))

    def dropout_mask(self, site, rows):
        """The inverted-dropout mask (1/(1-rate) or 0) of dropout site `site` for the CURRENT iterations value, as a
        device fp32 [rows, embedding_size] tensor (site numbering: include/b200xf.h).  For tests."""
        self._ensure_built()
        d = self.hparams["embedding_size"]
        out = torch.empty((rows, d), device=self.device, dtype=torch.float32)
        check(_lib.load().b200xf_dropout_mask(self._ctx, int(site), int(rows), d, ptr(out)))
        return out

    def synchronize(self):
        if self._ctx is not None:
            check(_lib.load().b200xf_synchronize(self._ctx))

    # ---- data parallel -----------------------------------------------------------------------------------
    def comm_init(self, unique_id: bytes, rank: int, world: int):
        self._ensure_built()
        buf = C.create_string_buffer(unique_id, 128)
        check(_lib.load().b200xf_comm_init(self._ctx, buf, int(rank), int(world)))
        self._comm = (rank, world)

    # ---- checkpoints (checkpoint.py:67,86; __main__.py:87) ------------------------------------------
    def save_weights(self, prefix):
        from .checkpoint_io import save_weights
        save_weights(self, prefix)

    def load_weights(self, prefix):
        from .checkpoint_io import load_weights
        load_weights(self, prefix)

    def save(self, path):
        from pathlib import Path
        Path(path).mkdir(parents=True, exist_ok=True)
        self.save_weights(str(Path(path) / "variables" / "variables"))


def greedy_decode(transformer, encoder_input, start_id, end_id, max_length=20, use_cache=True):
    """The decode loop of tester.py:13-53 without its tokenizers: starting from `start_id`, append the argmax of the last
    position, stop at `end_id` or after `max_length` tokens.  Returns (tokens [1, n] int64 on the device, attention_weights
    of the final pass over output[:, :-1]).

    The reference re-runs encoder and decoder over the whole prefix for every token (tester.py:32); with use_cache=True
    (default) the encoder output and the cross-attention K|V are computed once and every decoder layer keeps the K|V of the
    tokens decoded so far (b200xf_decode_begin / b200xf_decode_step) — the same tokens, one decoder position per step.
    use_cache=False is the reference's loop."""
    enc = _cuda_tokens(encoder_input, transformer.device)
    if enc.dim() == 1:
        enc = enc[None, :]
    assert enc.shape[0] == 1, "greedy_decode handles one sentence, as the reference's Tester"
    out = torch.full((1, 1), int(start_id), device=transformer.device, dtype=torch.int64)
    V = transformer.hparams["target_vocabulary_size"]
    nxt = torch.empty(1, device=transformer.device, dtype=torch.int64)
    lib = _lib.load()
    if use_cache:
        S = enc.shape[1]
        transformer._ensure_built(max(S, int(max_length) + 1))
        transformer._sync_weights()
        transformer._forward_serial += 1
        check(lib.b200xf_decode_begin(transformer._ctx, ptr(enc), 1, S, int(max_length) + 1))
        cur = out[:, 0].contiguous()
        for t in range(int(max_length)):
            check(lib.b200xf_decode_step(transformer._ctx, ptr(cur), t, None, ptr(nxt)))
            out = torch.cat([out, nxt[None, :]], dim=1)
            if int(nxt.item()) == int(end_id):
                break
            cur = nxt.clone()
    else:
        for _ in range(int(max_length)):
            logits, _ = transformer([enc, out], training=False)
            last = logits[0, -1]
            check(lib.b200xf_argmax_rows(ptr(last), V, 1, V, ptr(nxt), _lib.stream_ptr()))
            out = torch.cat([out, nxt[None, :]], dim=1)
            if int(nxt.item()) == int(end_id):
                break
    _, weights = transformer([enc, out[:, :-1]], training=False)
    return out, weights


def greedy_decode_batch(transformer, encoder_input, start_id, end_id, max_length=20):
    """Batched incremental greedy decoding over the K|V caches: every sentence of `encoder_input` [B, S] advances one
    position per step; a sentence that has produced `end_id` keeps emitting PAD (0).  Returns tokens [B, n] (device)."""
    enc = _cuda_tokens(encoder_input, transformer.device)
    B, S = enc.shape
    lib = _lib.load()
    transformer._ensure_built(max(B * S, B * (int(max_length) + 1)))
    transformer._sync_weights()
    transformer._forward_serial += 1
    check(lib.b200xf_decode_begin(transformer._ctx, ptr(enc), B, S, int(max_length) + 1))
    cur = torch.full((B,), int(start_id), device=transformer.device, dtype=torch.int64)
    done = torch.zeros(B, device=transformer.device, dtype=torch.bool)
    cols = [cur.clone()]
    nxt = torch.empty(B, device=transformer.device, dtype=torch.int64)
    for t in range(int(max_length)):
        check(lib.b200xf_decode_step(transformer._ctx, ptr(cur), t, None, ptr(nxt)))
        tok = torch.where(done, torch.zeros_like(nxt), nxt)
        cols.append(tok)
        done = done | (tok == int(end_id))
        if bool(done.all()):
            break
        cur = tok.clone()
    return torch.stack(cols, dim=1)


def build_model(transformer):
    """transformer_model.py:366-371 — creates the variables (the reference pushes zeros[64,1] through the model)."""
    transformer._ensure_built(64)
    z = torch.zeros((64, 1), dtype=torch.int64)
    transformer([z, z])


# ------------------------------------------------------------------ loss / accuracy (transformer_model.py:7-26)
def _ce_stats(actual, predicted, workers_count):
    """[loss, accuracy numerator, denominator] from fp32 logits, in fp32 (the reference computes both functions from the
    fp32 logits of Transformer.call; nothing is rounded to bf16 here)."""
    predicted = torch.as_tensor(predicted)
    if not predicted.is_cuda:
        raise _lib.B200XFError("local_loss_function/accuracy_function need device logits: there is no CPU fallback")
    actual = _cuda_tokens(actual, predicted.device).reshape(-1)
    V = predicted.shape[-1]
    rows = actual.numel()
    logits = predicted.reshape(rows, V).to(torch.float32).contiguous()
    out = torch.empty(3, device=predicted.device, dtype=torch.float32)
    scratch = torch.empty(2 * rows + 64, device=predicted.device, dtype=torch.float32)
    with torch.cuda.device(predicted.device):
        check(_lib.load().b200xf_vocab_ce_f32(ptr(logits), V, ptr(actual), rows, V, int(workers_count), ptr(out), ptr(scratch), _lib.stream_ptr()))
    return out


def local_loss_function(actual, predicted, workers_count):
    """transformer_model.py:11-17 — PAD-masked sparse softmax CE, / local non-pad count / workers_count."""
    return _ce_stats(actual, predicted, int(workers_count))[0]


def accuracy_function(actual, predicted):
    """transformer_model.py:20-26 — masked argmax accuracy."""
    s = _ce_stats(actual, predicted, 1)
    return s[1] / s[2]


loss_function = local_loss_function  # TF-tutorial name (BASELINE.json)
