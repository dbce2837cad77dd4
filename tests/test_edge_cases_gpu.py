"""Edge cases the reference's semantics imply: tiny shapes, build_model's [64,1] dummy batch, fully
padded sentences, padding inside a sentence, a batch whose targets are all PAD, out-of-range ids, the longest sequence
the positional table allows."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import b200xf
from oracle import transformer_oracle as O

HP = dict(layers_count=2, embedding_size=64, heads_count=4, feed_forward_hidden_size=128, input_vocabulary_size=203,
          target_vocabulary_size=157, maximum_input_length=200, maximum_target_length=200, droput_rate=0.0)


def _pair(seed=0):
    model = b200xf.Transformer(**HP, seed=seed)
    return model, O.OracleTransformer(**HP).load(model.get_weights_dict())


def _rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


def test_build_model_runs_the_reference_dummy_batch():
    model = b200xf.Transformer(**HP)
    b200xf.build_model(model)                                     # transformer_model.py:366-371: zeros [64, 1]
    assert len(model.trainable_variables) == 4 + 2 * (2 * 2) + 2 * (2 * 3) + 2 * 12 + 2 * 20
    logits, _ = model([np.zeros((64, 1), np.int64), np.zeros((64, 1), np.int64)])
    assert tuple(logits.shape) == (64, 1, 157) and torch.isfinite(logits).all()


@pytest.mark.parametrize("B,S,T", [(1, 1, 1), (1, 7, 1), (3, 1, 9), (2, 129, 130), (1, 200, 199)])
def test_small_and_maximum_shapes_match_oracle(B, S, T):
    model, oracle = _pair()
    inp, tgt = b200xf.synthetic_batch(B, S, T + 1, 203, 157, seed=S + T, ragged=False)
    logits, _ = model([inp, tgt[:, :-1]])
    ref, _ = oracle([torch.tensor(inp), torch.tensor(tgt[:, :-1])])
    assert _rel(logits.cpu().numpy(), ref.detach().numpy()) < 2e-2
    out = model.train_step_device(inp, tgt, 1).cpu().numpy()
    loss_ref, _, _ = O.train_step(oracle, inp, tgt, 1)
    assert abs(out[0] - loss_ref) < 1e-2


def test_fully_padded_and_interior_padding_sentences():
    """A source sentence that is entirely PAD makes every encoder / cross-attention row fully masked (uniform weights in
    the reference, not NaN); PAD in the middle of a sentence is masked by value, not by length (transformer_model.py:60)."""
    model, oracle = _pair(seed=2)
    inp, tgt = b200xf.synthetic_batch(4, 12, 11, 203, 157, seed=9, ragged=True)
    inp[1, :] = 0                      # whole source sentence is padding
    inp[2, 3] = 0; inp[2, 5] = 0       # interior padding
    tgt[3, 0] = 0                      # target starts with PAD: causal rows with no visible key
    tgt[0, 4] = 0                      # interior PAD label: masked in the loss
    logits, w = model([inp, tgt[:, :-1]])
    ref, ref_w = oracle([torch.tensor(inp), torch.tensor(tgt[:, :-1])])
    assert torch.isfinite(logits).all()
    assert _rel(logits.cpu().numpy(), ref.detach().numpy()) < 2e-2
    got = w["decoder_layer1_block2"].cpu()
    np.testing.assert_allclose(got[1].numpy(), 1.0 / 12, atol=1e-3)          # uniform over the 12 (all masked) keys
    assert float((got - ref_w["decoder_layer1_block2"].detach()).abs().max()) < 2e-2
    out = model.forward_backward_device(inp, tgt, 1).cpu().numpy()
    loss_ref, _, gref = O.train_step(oracle, inp, tgt, 1)
    assert abs(out[0] - loss_ref) < 1e-2
    grads = model.get_slot_dict("grad")
    assert all(np.isfinite(g).all() for g in grads.values())
    names = [n for n in gref if not n.endswith("key_generator_weights/bias")]
    cat = lambda g: np.concatenate([g[n].ravel() for n in names])
    assert _rel(cat(grads), cat(gref)) < 6e-2


def test_all_pad_targets_give_nan_loss_like_the_reference():
    model, _ = _pair()
    inp, tgt = b200xf.synthetic_batch(2, 6, 6, 203, 157, seed=1)
    tgt[:, 1:] = 0                     # no non-PAD label: sum(mask) == 0 -> 0/0 (transformer_model.py:16, unguarded)
    out = model.eval_step_device(inp, tgt).cpu().numpy()
    assert np.isnan(out[0]) and out[2] == 0


def test_out_of_range_token_ids_do_not_fault():
    model, _ = _pair()
    inp, tgt = b200xf.synthetic_batch(2, 6, 6, 203, 157, seed=1)
    inp[0, 2] = 100000; tgt[1, 2] = -5
    out = model.forward_backward_device(inp, tgt, 1).cpu().numpy()
    assert out.shape == (6,)
    torch.cuda.synchronize()
    logits, _ = model([inp, tgt[:, :-1]])
    assert tuple(logits.shape) == (2, 5, 157)


def test_sequence_longer_than_the_positional_table_is_rejected():
    model, _ = _pair()
    with pytest.raises(b200xf.B200XFError, match="maximum_input_length"):
        model([np.ones((1, 201), np.int64), np.ones((1, 3), np.int64)])
