"""Forward-only API: the attention-weight maps Transformer.call returns (transformer_model.py:312-313)
and the greedy decode loop of tester.py:13-53, against the CPU oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import b200xf
from oracle import transformer_oracle as O

HP = dict(layers_count=2, embedding_size=64, heads_count=4, feed_forward_hidden_size=128, input_vocabulary_size=203,
          target_vocabulary_size=157, maximum_input_length=64, maximum_target_length=64, droput_rate=0.1)


def test_attention_weight_maps_match_oracle():
    model = b200xf.Transformer(**HP, seed=6)
    oracle = O.OracleTransformer(**HP).load(model.get_weights_dict())
    inp, tgt = b200xf.synthetic_batch(3, 22, 20, 203, 157, seed=2, ragged=True)
    logits, weights = model([inp, tgt[:, :-1]], training=False)
    ref_logits, ref_w = oracle([torch.tensor(inp), torch.tensor(tgt[:, :-1])])
    assert sorted(weights.keys()) == sorted(ref_w.keys()) == ["decoder_layer1_block1", "decoder_layer1_block2",
                                                                "decoder_layer2_block1", "decoder_layer2_block2"]
    for k, ref in ref_w.items():
        got = weights[k].cpu()
        assert tuple(got.shape) == tuple(ref.shape)
        assert float((got - ref.detach()).abs().max()) < 2e-2, k             # probabilities; bf16 Q/K
        np.testing.assert_allclose(got.sum(-1).numpy(), 1.0, atol=2e-3)        # rows are distributions
    # look-ahead mask: nothing above the diagonal; padded source keys get zero weight
    w1 = weights["decoder_layer1_block1"].cpu()
    assert float(torch.triu(w1, diagonal=1).abs().max()) == 0.0
    w2 = weights["decoder_layer2_block2"].cpu()
    pad = torch.tensor(inp == 0)[:, None, None, :].expand_as(w2)
    assert float(w2[pad].abs().max()) == 0.0
    _, stale = model([inp, tgt[:, :-1]], training=False)
    model([inp, tgt[:, :-1]], training=False)
    with pytest.raises(b200xf.B200XFError):
        stale["decoder_layer2_block1"]                                         # the model has run another batch since


def test_greedy_decode_follows_the_oracle():
    model = b200xf.Transformer(**HP, seed=8)
    w = model.get_weights_dict()
    w["final_layer/kernel"] = w["final_layer/kernel"] * 8.0                    # sharpen the logits of the random model
    model.set_weights_dict(w)
    oracle = O.OracleTransformer(**HP).load(w)
    src, _ = b200xf.synthetic_batch(1, 12, 5, 203, 157, seed=4, ragged=False)
    tokens, weights = b200xf.greedy_decode(model, src, start_id=2, end_id=3, max_length=12)
    toks = tokens.cpu().numpy()[0]
    assert toks[0] == 2 and 2 <= len(toks) <= 13
    assert (toks[1:-1] != 3).all()                                             # stops at the first END
    for n in range(1, len(toks)):                                              # every chosen token is the oracle's argmax (up to bf16 near-ties)
        ref, _ = oracle([torch.tensor(src), torch.tensor(toks[None, :n])])
        last = ref[0, -1].detach()
        assert float(last.max() - last[toks[n]]) < 5e-2, (n, int(last.argmax()), int(toks[n]))
    assert tuple(weights["decoder_layer1_block2"].shape) == (1, 4, len(toks) - 1, 12)


def test_argmax_rows_first_max_wins():
    x = torch.tensor([[1.0, 5.0, 5.0, 2.0], [7.0, 7.0, 7.0, 7.0], [-1.0, -3.0, -0.5, -0.5]], device="cuda")
    out = torch.empty(3, device="cuda", dtype=torch.int64)
    b200xf._lib.check(b200xf._lib.load().b200xf_argmax_rows(b200xf._lib.ptr(x), 4, 3, 4, b200xf._lib.ptr(out), b200xf._lib.stream_ptr()))
    assert out.tolist() == [1, 0, 2]


BASEHP = dict(layers_count=2, embedding_size=512, heads_count=8, feed_forward_hidden_size=1024, input_vocabulary_size=3000,
              target_vocabulary_size=3000, maximum_input_length=256, maximum_target_length=256, droput_rate=0.0)


@pytest.mark.parametrize("hp,S,steps", [(HP, 12, 14), (BASEHP, 40, 24)])
def test_kv_cache_decode_equals_the_reencode_loop_and_the_oracle(hp, S, steps):
    """greedy_decode with the K|V caches (encoder and cross-attention K|V once, one decoder position per step) emits the same
    tokens as the reference's loop that re-runs the whole model per token (tester.py:27-42), and each step's logits equal the
    last row of the full forward pass over the prefix; the oracle agrees with every chosen token."""
    vin, vt = hp["input_vocabulary_size"], hp["target_vocabulary_size"]
    model = b200xf.Transformer(**hp, seed=8)
    w = model.get_weights_dict()
    w["final_layer/kernel"] = w["final_layer/kernel"] * 8.0
    model.set_weights_dict(w)
    oracle = O.OracleTransformer(**hp).load(w)
    src, _ = b200xf.synthetic_batch(1, S, 5, vin, vt, seed=4, ragged=False)
    cached, _ = b200xf.greedy_decode(model, src, start_id=2, end_id=3, max_length=steps, use_cache=True)
    plain, _ = b200xf.greedy_decode(model, src, start_id=2, end_id=3, max_length=steps, use_cache=False)
    assert cached.tolist() == plain.tolist()
    toks = cached.cpu().numpy()[0]
    for n in range(1, len(toks)):
        ref, _ = oracle([torch.tensor(src), torch.tensor(toks[None, :n])])
        last = ref[0, -1].detach()
        assert float(last.max() - last[toks[n]]) < 5e-2, (n, int(last.argmax()), int(toks[n]))
    # per-step logits against the full pass over the same prefix (same kernels, same accumulation order per row)
    lib = b200xf._lib.load(); ptr = b200xf._lib.ptr; check = b200xf._lib.check
    enc = torch.tensor(src, device="cuda")
    check(lib.b200xf_decode_begin(model._ctx, ptr(enc), 1, S, len(toks)))
    step_logits = torch.empty((1, vt), device="cuda")
    for t in range(len(toks) - 1):
        cur = torch.tensor([int(toks[t])], device="cuda")
        check(lib.b200xf_decode_step(model._ctx, ptr(cur), t, ptr(step_logits), None))
        full, _ = model([src, toks[None, :t + 1]], training=False)
        rel = float((step_logits[0] - full[0, -1]).norm() / full[0, -1].norm())
        assert rel < 2e-3, (t, rel)
        check(lib.b200xf_decode_begin(model._ctx, ptr(enc), 1, S, len(toks)))   # the full pass reused the caches' buffers: replay the prefix
        for u in range(t + 1):
            check(lib.b200xf_decode_step(model._ctx, ptr(torch.tensor([int(toks[u])], device="cuda")), u, None, None))


def test_batched_kv_cache_decode_matches_single_sentence_decodes():
    model = b200xf.Transformer(**HP, seed=8)
    w = model.get_weights_dict()
    w["final_layer/kernel"] = w["final_layer/kernel"] * 8.0
    model.set_weights_dict(w)
    src, _ = b200xf.synthetic_batch(5, 12, 5, 203, 157, seed=14, ragged=True)
    batch = b200xf.greedy_decode_batch(model, src, start_id=2, end_id=3, max_length=12).cpu().numpy()
    for b in range(5):
        one, _ = b200xf.greedy_decode(model, src[b:b + 1], start_id=2, end_id=3, max_length=12)
        one = one.cpu().numpy()[0]
        assert batch[b, :len(one)].tolist() == one.tolist(), b
        assert (batch[b, len(one):] == 0).all()                                # PAD after END


def test_loss_functions_use_fp32_logits():
    """local_loss_function / accuracy_function (transformer_model.py:11-26) on the fp32 logits Transformer.call returns: the loss
    equals torch's fp32 cross-entropy of the SAME logits to fp32 accuracy, and a near-tie that bf16 rounding would flip is
    resolved as in fp32."""
    g = torch.Generator(device="cuda").manual_seed(3)
    logits = torch.randn(4, 9, 157, device="cuda", generator=g) * 3
    labels = torch.randint(1, 157, (4, 9), device="cuda", generator=g)
    labels[0, -3:] = 0
    logits[1, 2, 40] = 20.0; logits[1, 2, 41] = 20.0 + 2e-3; labels[1, 2] = 41  # equal after bf16 rounding (argmax would be 40)
    l = float(b200xf.local_loss_function(labels, logits, 2)); acc = float(b200xf.accuracy_function(labels, logits))
    mask = labels != 0
    ce = torch.nn.functional.cross_entropy(logits.reshape(-1, 157), labels.reshape(-1), reduction="none").reshape(4, 9)
    ref_l = float((ce * mask).sum() / mask.sum() / 2)
    ref_acc = float(((logits.argmax(-1) == labels) & mask).sum() / mask.sum())
    assert abs(l - ref_l) < 2e-5 * max(1.0, abs(ref_l)), (l, ref_l)
    assert acc == pytest.approx(ref_acc, abs=1e-7)
