"""Known-answer tests that pin the oracle to facts derivable from the reference source
 and to the checkpoint index fixture decoded from the reference artefact."""
import json
from pathlib import Path

import numpy as np
import torch

from oracle import transformer_oracle as O

GOLD = json.loads((Path(__file__).parent / "golden" / "param_table.json").read_text())
TUT = dict(layers_count=4, embedding_size=128, heads_count=8, feed_forward_hidden_size=512,
           input_vocabulary_size=7765, target_vocabulary_size=7010,
           maximum_input_length=1000, maximum_target_length=1000, droput_rate=0.0)


def test_look_ahead_mask_kat():  # transformer_model.py:65-70
    assert O.create_look_ahead_mask(3).tolist() == [[0, 1, 1], [0, 0, 1], [0, 0, 0]]


def test_padding_mask_kat():  # transformer_model.py:56-62
    m = O.create_padding_mask(torch.tensor([[7, 6, 0, 0, 1], [1, 2, 3, 0, 0], [0, 0, 0, 4, 5]]))
    assert tuple(m.shape) == (3, 1, 1, 5)
    assert m[:, 0, 0].tolist() == [[0, 0, 1, 1, 0], [0, 0, 0, 1, 1], [1, 1, 1, 0, 0]]


def test_combined_decoder_mask_kat():  # transformer_model.py:359-362
    _, la, _ = O.create_masks(torch.tensor([[1, 2, 3]]), torch.tensor([[5, 9, 0]]))
    assert la[0, 0].tolist() == [[0, 1, 1], [0, 0, 1], [0, 0, 1]]


def test_positional_encoding_kat():  # transformer_model.py:29-53
    pe = O.positional_encoding(1000, 128)
    assert tuple(pe.shape) == (1, 1000, 128)
    assert pe[0, 0].tolist() == [0.0, 1.0] * 64
    np.testing.assert_allclose(pe[0, 1, :4].numpy(), [0.841471, 0.5403023, 0.7617204, 0.6479058], rtol=2e-6)
    np.testing.assert_allclose(pe[0, 1, -2:].numpy(), [1.1547821e-4, 1.0], rtol=2e-6)


def test_learning_rate_kat():  # __main__.py:56-67
    assert O.learning_rate(0, 128) == 0.0
    for step, want in [(1, 3.4938563e-07), (100, 3.49385627e-05), (4000, 1.39754254e-03), (10000, 8.83883447e-04)]:
        np.testing.assert_allclose(O.learning_rate(step, 128), want, rtol=3e-6)


def test_adam_tf_form_vs_torch_form():  # a19: epsilon placement differs by ~0.6 % at g=1e-6, t=1
    g = np.array([1e-6], np.float32); z = np.zeros(1, np.float32)
    w, m, v = O.adam_update_numpy(z, g, z, z, iterations=0, lr=1.0)
    np.testing.assert_allclose(m, 0.1 * g, rtol=1e-6); np.testing.assert_allclose(v, 0.02 * g * g, rtol=1e-5)
    torch_form = -(0.1 * 1e-6 / 0.1) / (np.sqrt(0.02 * 1e-12 / 0.02) + 1e-9)
    rel = (w[0] - torch_form) / torch_form
    assert -0.0075 < rel < -0.0045, rel


def test_masked_row_is_uniform():  # fp32 (-1e9)+x == -1e9 for |x| < 32
    x = torch.tensor([[3.0, -7.5, 12.25]]) + torch.ones(1, 3) * O.NEG_MASK
    assert torch.softmax(x, -1).tolist() == [[1 / 3] * 3] or torch.allclose(torch.softmax(x, -1), torch.full((1, 3), 1 / 3))


def test_parameter_table_matches_reference_index():
    names = O.parameter_names(4)
    shapes = O.parameter_shapes(4, 128, 512, 7765, 7010)
    gold = {t["name"]: t for t in GOLD["tensors"]}
    assert len(names) == GOLD["n_tensors"] == 172
    assert set(names) == set(gold)
    # oracle enumerates in the data shard's physical order
    assert names == [t["name"] for t in GOLD["tensors"]]
    assert all(list(shapes[n]) == gold[n]["shape"] for n in names)
    assert sum(int(np.prod(s)) for s in shapes.values()) == GOLD["total_params"] == 4646882


def _tiny(seed=0, emulate=False):
    hp = dict(layers_count=2, embedding_size=32, heads_count=4, feed_forward_hidden_size=64,
              input_vocabulary_size=50, target_vocabulary_size=40, maximum_input_length=64,
              maximum_target_length=64, droput_rate=0.0)
    m = O.OracleTransformer(**hp, emulate_bf16=emulate)
    m.load(O.init_parameters(2, 32, 64, 50, 40, seed))
    return m


def test_first_step_leaves_weights_bit_identical():  # lr(0) == 0
    m = _tiny()
    before = {n: p.detach().numpy().copy() for n, p in m.params.items()}
    inp, tgt = O.synthetic_batch(4, 10, 11, 50, 40, seed=5)
    loss, logits, grads = O.train_step(m, inp, tgt)
    for n, p in m.params.items():
        assert np.array_equal(before[n], p.detach().numpy()), n
        np.testing.assert_allclose(m.m[n], 0.1 * grads[n], rtol=1e-6, atol=1e-30)
        np.testing.assert_allclose(m.v[n], 0.02 * grads[n] ** 2, rtol=1e-5, atol=1e-30)
    assert abs(loss - np.log(40)) < 0.5  # fresh-init loss ~ ln V


def test_loss_decreases_and_workers_scaling():
    m = _tiny(); m.iterations = 1000
    inp, tgt = O.synthetic_batch(4, 10, 11, 50, 40, seed=5)
    l0, *_ = O.train_step(m, inp, tgt)
    for _ in range(5):
        l1, *_ = O.train_step(m, inp, tgt)
    assert l1 < l0
    a = _tiny(); b = _tiny()
    la, _, ga = O.train_step(a, inp, tgt, workers_count=1)
    lb, _, gb = O.train_step(b, inp, tgt, workers_count=3)
    np.testing.assert_allclose(lb * 3, la, rtol=1e-6)
    n = "final_layer/kernel"
    np.testing.assert_allclose(gb[n] * 3, ga[n], rtol=1e-4, atol=1e-9)


def test_padded_positions_do_not_change_valid_logits():
    m = _tiny()
    inp, tgt = O.synthetic_batch(2, 8, 9, 50, 40, seed=3, ragged=False)
    lg, _ = m([torch.tensor(inp), torch.tensor(tgt[:, :-1])])
    inp2 = np.concatenate([inp, np.zeros((2, 3), np.int64)], 1)
    tgt2 = np.concatenate([tgt[:, :-1], np.zeros((2, 2), np.int64)], 1)
    lg2, _ = m([torch.tensor(inp2), torch.tensor(tgt2)])
    torch.testing.assert_close(lg2[:, :8], lg, rtol=1e-4, atol=1e-5)


def test_argmax_first_max_wins():
    x = torch.tensor([[[1.0, 5.0, 5.0, 2.0]]])
    num, den = O.accuracy_counts(torch.tensor([[1]]), x)
    assert (num, den) == (1, 1)
    num, den = O.accuracy_counts(torch.tensor([[2]]), x)
    assert (num, den) == (0, 1)


def test_sdpa_closed_form_kat():  # transformer_model.py:73-109 on operands whose softmax is (numerically) one-hot or an exact tie
    k = torch.tensor([[10., 0, 0], [0, 10, 0], [0, 0, 10], [0, 0, 10]])
    v = torch.tensor([[1., 0], [10, 0], [100, 5], [1000, 6]])
    m = O.OracleTransformer(1, 4, 1, 4, 8, 8, 8, 8)
    for q, w_want, o_want in (([[0., 10, 0]], [[0, 1, 0, 0]], [[10, 0]]),
                              ([[0., 0, 10]], [[0, 0, .5, .5]], [[550, 5.5]]),
                              ([[10., 10, 0]], [[.5, .5, 0, 0]], [[5.5, 0]])):
        o, w = m._sdpa(torch.tensor(q), k, v, None)
        np.testing.assert_allclose(w.numpy(), w_want, atol=1e-6)     # e^(-100/sqrt(3)) ~ 1e-25 off a one-hot
        np.testing.assert_allclose(o.numpy(), o_want, rtol=1e-6, atol=1e-20)
    # a masked key gets exactly zero weight in fp32 (additive -1e9, :101-102), whatever its score
    o, w = m._sdpa(torch.tensor([[0., 0, 10]]), k, v, torch.tensor([[0., 0, 0, 1]]))
    assert w[0, 3].item() == 0.0 and w[0, 2].item() == 1.0
    np.testing.assert_allclose(o.numpy(), [[100, 5]], rtol=1e-6)


def test_forward_matches_an_independent_implementation_of_the_architecture():
    """Not a pin to the reference (TensorFlow cannot run here): a cross-check of the restatement against torch.nn's own,
    independently written post-LayerNorm encoder / decoder stacks with the SAME weights -- head split and merge order, the
    1/sqrt(dk) scaling, where the residuals and LayerNorms sit, epsilon, biased variance, ReLU feed-forward, key-padding and
    look-ahead masking (transformer_model.py:112-166,178-248).  Embedding, sqrt(d) scaling and positional encoding are the
    oracle's own on both sides (torch.nn has none)."""
    import torch.nn as nn
    L, d, H, dff, V = 2, 32, 4, 64, 50
    model = O.OracleTransformer(L, d, H, dff, V, V, 64, 64).load(
        {n: a + (0.1 * np.random.default_rng(5).standard_normal(a.shape).astype(np.float32) if a.ndim == 1 else 0)
         for n, a in O.init_parameters(L, d, dff, V, V, seed=3).items()})   # non-trivial biases, gammas, betas
    P = {n: t.detach() for n, t in model.params.items()}
    enc = nn.TransformerEncoder(nn.TransformerEncoderLayer(d, H, dff, dropout=0.0, activation="relu", layer_norm_eps=1e-6,
                                                           batch_first=True, norm_first=False), L, enable_nested_tensor=False)
    dec = nn.TransformerDecoder(nn.TransformerDecoderLayer(d, H, dff, dropout=0.0, activation="relu", layer_norm_eps=1e-6,
                                                           batch_first=True, norm_first=False), L)

    def put_mha(mod, prefix):   # nn.MultiheadAttention packs q | k | v as rows of in_proj_weight [3d, d] (y = x W^T + b)
        g = lambda n: P[f"{prefix}/{n}"]
        mod.in_proj_weight.data = torch.cat([g("query_generator_weights/kernel").T, g("key_generator_weights/kernel").T,
                                             g("value_generator_weights/kernel").T])
        mod.in_proj_bias.data = torch.cat([g("query_generator_weights/bias"), g("key_generator_weights/bias"), g("value_generator_weights/bias")])
        mod.out_proj.weight.data = g("dense/kernel").T.contiguous(); mod.out_proj.bias.data = g("dense/bias").clone()

    def put(lin_or_ln, prefix, dense=True):
        if dense:
            lin_or_ln.weight.data = P[prefix + "/kernel"].T.contiguous(); lin_or_ln.bias.data = P[prefix + "/bias"].clone()
        else:
            lin_or_ln.weight.data = P[prefix + "/gamma"].clone(); lin_or_ln.bias.data = P[prefix + "/beta"].clone()

    for i, lyr in enumerate(enc.layers):
        p = f"encoder/encoder_layers/{i}"
        put_mha(lyr.self_attn, p + "/mha")
        put(lyr.linear1, p + "/ffn/layer_with_weights-0"); put(lyr.linear2, p + "/ffn/layer_with_weights-1")
        put(lyr.norm1, p + "/layernorm1", False); put(lyr.norm2, p + "/layernorm2", False)
    for i, lyr in enumerate(dec.layers):
        p = f"decoder/decoder_layers/{i}"
        put_mha(lyr.self_attn, p + "/mha1"); put_mha(lyr.multihead_attn, p + "/mha2")
        put(lyr.linear1, p + "/ffn/layer_with_weights-0"); put(lyr.linear2, p + "/ffn/layer_with_weights-1")
        for k in (1, 2, 3):
            put(getattr(lyr, f"norm{k}"), p + f"/layernorm{k}", False)
    enc.train(); dec.train()    # dropout is 0; training mode keeps torch on its plain (non-fused) code path

    inp, tgt = (torch.as_tensor(a) for a in O.synthetic_batch(3, 11, 10, V, V, seed=2, ragged=True))
    tgt = tgt[:, :-1]
    with torch.no_grad():
        want, _ = model([inp, tgt])
        emb = lambda name, tok, pe: P[name][tok] * np.sqrt(np.float32(d)) + pe[:, :tok.shape[1], :]
        src_pad, tgt_pad = inp == 0, tgt == 0
        memory = enc(emb("encoder/embedding/embeddings", inp, model.pe_in), src_key_padding_mask=src_pad)
        causal = torch.triu(torch.ones(tgt.shape[1], tgt.shape[1], dtype=torch.bool), 1)
        out = dec(emb("decoder/embedding/embeddings", tgt, model.pe_tgt), memory, tgt_mask=causal, tgt_key_padding_mask=tgt_pad,
                  memory_key_padding_mask=src_pad)
        got = out @ P["final_layer/kernel"] + P["final_layer/bias"]
    valid = ~tgt_pad                                   # rows whose query is PAD are never used by the loss (transformer_model.py:11-17)
    assert valid.sum() > 10 and src_pad.any() and tgt_pad.any()
    np.testing.assert_allclose(got[valid].numpy(), want[valid].numpy(), rtol=2e-4, atol=2e-5)
