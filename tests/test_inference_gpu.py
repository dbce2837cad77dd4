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
