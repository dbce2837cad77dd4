"""End-to-end parity of the CUDA train step against the CPU fp32 oracle (oracle/transformer_oracle.py) on the same
seeded synthetic batches and the same initial weights, dropout 0.  The CUDA path computes in bf16 with fp32
accumulation and fp32 master weights; tolerances are written next to each assertion."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import b200xf
from oracle import transformer_oracle as O

SMALL = dict(layers_count=2, embedding_size=64, heads_count=4, feed_forward_hidden_size=128, input_vocabulary_size=203,
             target_vocabulary_size=157, maximum_input_length=64, maximum_target_length=64, droput_rate=0.0)
TUT = dict(layers_count=4, embedding_size=128, heads_count=8, feed_forward_hidden_size=512, input_vocabulary_size=7765,
           target_vocabulary_size=7010, maximum_input_length=1000, maximum_target_length=1000, droput_rate=0.0)


def _pair(hp, seed=0, emulate_bf16=False):
    model = b200xf.Transformer(**hp, seed=seed)
    weights = model.get_weights_dict()
    oracle = O.OracleTransformer(**hp, emulate_bf16=emulate_bf16).load(weights)
    return model, oracle


def _rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


def test_parameter_table_matches_reference_checkpoint_index():
    import json
    from pathlib import Path
    gold = json.loads((Path(__file__).parent / "golden" / "param_table.json").read_text())
    model = b200xf.Transformer(**TUT)
    got = [(v.name, list(v.shape)) for v in model.trainable_variables]
    assert got == [(t["name"], t["shape"]) for t in gold["tensors"]]
    assert sum(int(np.prod(s)) for _, s in got) == gold["total_params"] == 4646882


@pytest.mark.parametrize("hp,B,S,T", [(SMALL, 4, 20, 17), (TUT, 64, 40, 40)])
def test_forward_logits_parity(hp, B, S, T):
    model, oracle = _pair(hp)
    inp, tgt = b200xf.synthetic_batch(B, S, T + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=1234, ragged=True)
    logits, _ = model([inp, tgt[:, :-1]], training=False)
    ref, _ = oracle([torch.tensor(inp), torch.tensor(tgt[:, :-1])])
    got = logits.cpu().numpy(); ref = ref.detach().numpy()
    assert np.isfinite(got).all()
    assert _rel(got, ref) < 2e-2, _rel(got, ref)  # logits rel-L2 <= 2e-2
    # against the oracle with bf16-rounded GEMM operands the gap closes further (rounding, not logic)
    _, oracle16 = _pair(hp, emulate_bf16=True)
    ref16, _ = oracle16([torch.tensor(inp), torch.tensor(tgt[:, :-1])])
    assert _rel(got, ref16.detach().numpy()) < 1.5e-2


@pytest.mark.parametrize("hp,B,S,T", [(SMALL, 4, 20, 17), (TUT, 64, 40, 40)])
def test_loss_accuracy_and_gradients_parity(hp, B, S, T):
    model, oracle = _pair(hp)
    _, oracle16 = _pair(hp, emulate_bf16=True)
    inp, tgt = b200xf.synthetic_batch(B, S, T + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=77, ragged=True)
    out = model.forward_backward_device(inp, tgt, workers_count=3).cpu().numpy()
    loss_ref, logits_ref, grads_ref = O.train_step(oracle, inp, tgt, workers_count=3)
    _, _, grads16 = O.train_step(oracle16, inp, tgt, workers_count=3)
    assert abs(out[0] - loss_ref) < 1e-2 / 3, (out[0], loss_ref)            # loss |delta| <= 1e-2 (here / workers)
    num, den = O.accuracy_counts(torch.tensor(tgt[:, 1:]), logits_ref)
    assert int(out[2]) == den                                               # PAD mask count is exact
    assert abs(int(out[1]) - num) <= max(2, den // 200)                     # argmax flips only on bf16 near-ties
    grads = model.get_slot_dict("grad")
    worst = ("", 0.0)
    for name, gref in grads_ref.items():
        if name.endswith("key_generator_weights/bias"):
            # softmax is invariant to a per-query constant, so d(loss)/d(key bias) is exactly 0 in real arithmetic:
            # the oracle holds fp32 noise, the CUDA path bf16 noise; require it to be small next to the query-bias gradient
            qn = np.linalg.norm(grads_ref[name.replace("key_generator", "query_generator")])
            assert np.linalg.norm(grads[name]) < 0.25 * qn + 1e-7, (name, np.linalg.norm(grads[name]), qn)
            continue
        if np.linalg.norm(gref) < 1e-12:
            assert np.linalg.norm(grads[name]) < 1e-6, name
            continue
        e = _rel(grads[name], gref)
        if e > worst[1]: worst = (name, e)
    # Measured (scripts/parity_report.py): whole-gradient rel-L2 1.75e-2 on the tutorial config and 4.0e-2 on the
    # 68-token toy config, against 1.69e-2 / 3.4e-2 for the fp32 oracle run with bf16-rounded GEMM operands: the gap
    # is bf16 rounding.  The worst single tensor (top decoder layer's self-attention query kernel, whose gradient is
    # the small difference P*(dP-delta)) reaches 0.15.
    assert worst[1] < 0.2, worst                                            # per-tensor gradient rel-L2
    names = [n for n in grads_ref if not n.endswith("key_generator_weights/bias")]
    cat = lambda g: np.concatenate([g[n].ravel() for n in names])
    e, e16 = _rel(cat(grads), cat(grads_ref)), _rel(cat(grads16), cat(grads_ref))
    assert e < 5e-2 and e < 1.3 * e16 + 5e-3, (e, e16)                      # whole-gradient rel-L2, and vs bf16 rounding alone


@pytest.mark.parametrize("hp,B,S,T", [(SMALL, 4, 20, 17), (TUT, 64, 40, 40)])
def test_train_steps_parity(hp, B, S, T):
    model, oracle = _pair(hp)
    oracle16 = O.OracleTransformer(**hp, emulate_bf16=True).load(model.get_weights_dict())
    vin, vt = hp["input_vocabulary_size"], hp["target_vocabulary_size"]
    w0 = model.get_weights_dict()
    # step 0: lr == 0 -> weights bit-identical, m = 0.1 g, v = 0.02 g^2
    inp, tgt = b200xf.synthetic_batch(B, S, T + 1, vin, vt, seed=1234, ragged=True)
    model.train_step_device(inp, tgt, 1); O.train_step(oracle, inp, tgt, 1); O.train_step(oracle16, inp, tgt, 1)
    w1 = model.get_weights_dict()
    assert all(np.array_equal(w0[k], w1[k]) for k in w0)
    assert model.iterations == 1
    # continue from a warmed-up iteration count so that lr is material
    model.iterations = 2000; oracle.iterations = 2000; oracle16.iterations = 2000
    losses = []
    for step in range(4):
        inp, tgt = b200xf.synthetic_batch(B, S, T + 1, vin, vt, seed=1235 + step, ragged=True)
        res = model.train_step_device(inp, tgt, 1).cpu().numpy()
        ref_loss, _, gref = O.train_step(oracle, inp, tgt, 1)
        O.train_step(oracle16, inp, tgt, 1)
        losses.append((res[0], ref_loss))
        assert abs(res[0] - ref_loss) < 2e-2, (step, res[0], ref_loss)      # loss curve within 2e-2 at every step
    assert model.iterations == 2004
    def delta_w_error(other):
        num = den = 0.0; agree = tot = 0.0
        for name, p in oracle.params.items():
            ref = p.detach().numpy(); d_ref = ref - w1[name]; d_got = other[name] - w1[name]
            sel = np.abs(gref[name]) > 1e-7
            if sel.any():
                num += float(np.sum((d_got[sel] - d_ref[sel]) ** 2)); den += float(np.sum(d_ref[sel] ** 2))
                wgt = np.abs(gref[name][sel])
                agree += float(np.sum(wgt * (np.sign(d_got[sel]) == np.sign(d_ref[sel])))); tot += float(np.sum(wgt))
        return np.sqrt(num / den), agree / tot

    err, sign = delta_w_error(model.get_weights_dict())
    err16, sign16 = delta_w_error({k: p.detach().numpy() for k, p in oracle16.params.items()})
    # Early Adam is sign-like (dw ~ lr * g / |g|), so elements whose gradient is within bf16 noise move by O(lr) either
    # way: measured 0.11 on the tutorial config, against 0.10 for the fp32 oracle with bf16-rounded GEMM operands.
    assert err < 0.15, err                                                  # delta-w rel-L2 where |g_ref| > 1e-7
    assert err < 1.3 * err16 + 0.01, (err, err16)                           # no worse than bf16 rounding alone explains
    # |g|-weighted sign agreement: 0.993 on the tutorial config (0.988 on the 68-token toy config, where the bf16-rounded
    # oracle itself only reaches 0.986)
    assert sign > min(0.99, sign16 - 0.005), (sign, sign16)


def test_host_entry_point_matches_device_entry_point():
    a = b200xf.Transformer(**SMALL, seed=3); b = b200xf.Transformer(**SMALL, seed=3)
    a.iterations = 500; b.iterations = 500
    inp, tgt = b200xf.synthetic_batch(4, 20, 18, 203, 157, seed=5, ragged=True)
    ra = a.train_step_device(inp, tgt, 1).cpu().numpy()
    rb = np.array(b.train_step_host(torch.tensor(inp).pin_memory(), torch.tensor(tgt).pin_memory(), 1))
    np.testing.assert_allclose(ra[:3], rb[:3], rtol=1e-5)
    wa, wb = a.get_weights_dict(), b.get_weights_dict()
    assert max(_rel(wa[k], wb[k]) for k in wa) < 1e-5  # identical up to split-K atomic ordering


def test_eval_step_and_loss_functions():
    model, oracle = _pair(SMALL)
    inp, tgt = b200xf.synthetic_batch(4, 20, 18, 203, 157, seed=9, ragged=True)
    res = model.eval_step_device(inp, tgt).cpu().numpy()
    logits, _ = model([inp, tgt[:, :-1]])
    l = b200xf.local_loss_function(tgt[:, 1:], logits, 1); acc = b200xf.accuracy_function(tgt[:, 1:], logits)
    ref_logits, _ = oracle([torch.tensor(inp), torch.tensor(tgt[:, :-1])])
    ref_l = float(O.local_loss_function(torch.tensor(tgt[:, 1:]), ref_logits, 1))
    assert abs(res[0] - ref_l) < 1e-2 and abs(float(l) - ref_l) < 1e-2
    assert abs(float(acc) - res[1] / res[2]) < 0.02
    w0 = model.get_weights_dict(); model.eval_step_device(inp, tgt); w1 = model.get_weights_dict()
    assert all(np.array_equal(w0[k], w1[k]) for k in w0)


def test_variable_shapes_and_workspace_growth():
    model = b200xf.Transformer(**SMALL, max_tokens=64)
    w0 = model.get_weights_dict()
    for (B, S, T1) in [(2, 8, 9), (4, 30, 25), (3, 17, 40)]:
        inp, tgt = b200xf.synthetic_batch(B, S, T1, 203, 157, seed=B, ragged=True)
        out = model.train_step_device(inp, tgt, 1).cpu().numpy()
        assert np.isfinite(out).all()
    assert model.iterations == 3
    with pytest.raises(b200xf.B200XFError):
        model([np.ones((1, 100), np.int64), np.ones((1, 4), np.int64)])  # longer than maximum_input_length


def test_dropout_train_step_parity_with_injected_masks():
    """droput_rate = 0.1 (the reference script's value, __main__.py:43): the CUDA path's masks are read back through
    b200xf_dropout_mask and injected into the oracle, so forward, loss and every gradient can be compared exactly as
    in the rate-0 tests; plus the statistics of the mask itself."""
    hp = dict(SMALL, droput_rate=0.1)
    model = b200xf.Transformer(**hp, seed=4)
    oracle = O.OracleTransformer(**hp).load(model.get_weights_dict())
    oracle16 = O.OracleTransformer(**hp, emulate_bf16=True).load(model.get_weights_dict())
    B, S, T = 6, 24, 21
    inp, tgt = b200xf.synthetic_batch(B, S, T + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=31, ragged=True)
    model.iterations = 700; oracle.iterations = 700; oracle16.iterations = 700
    L = hp["layers_count"]
    masks = {}
    for site in range(2 + 2 * L + 3 * L):
        rows = B * S if (site == 0 or 2 <= site < 2 + 2 * L) else B * T
        masks[site] = model.dropout_mask(site, rows).cpu()
    keep = torch.cat([m.reshape(-1) for m in masks.values()])
    assert set(torch.unique(keep).tolist()) == {0.0, float(np.float32(1.0) / (np.float32(1.0) - np.float32(0.1)))}
    frac = float((keep == 0).float().mean())
    assert abs(frac - 0.1) < 0.01, frac                                      # ~1e5 Bernoulli(0.1) draws: sigma ~ 1e-3
    assert not torch.equal(masks[2], masks[3])                                # sites draw independent masks
    oracle.dropout_masks = masks; oracle16.dropout_masks = masks
    out = model.forward_backward_device(inp, tgt, 1).cpu().numpy()
    grads = model.get_slot_dict("grad")
    loss_ref, _, gref = O.train_step(oracle, inp, tgt, 1)
    _, _, g16 = O.train_step(oracle16, inp, tgt, 1)
    assert abs(out[0] - loss_ref) < 1e-2, (out[0], loss_ref)
    names = [n for n in gref if not n.endswith("key_generator_weights/bias")]
    cat = lambda g: np.concatenate([g[n].ravel() for n in names])
    e, e16 = _rel(cat(grads), cat(gref)), _rel(cat(g16), cat(gref))
    assert e < 6e-2 and e < 1.3 * e16 + 5e-3, (e, e16)                        # same rule as the rate-0 gradient test
    # a different step draws a different mask; the same step reproduces it
    m_again = model.dropout_mask(2, B * S).cpu()
    assert torch.equal(m_again, masks[2])
    model.iterations = 701
    assert not torch.equal(model.dropout_mask(2, B * S).cpu(), masks[2])
    # training=False ignores dropout (fresh oracle: the one above has meanwhile applied its Adam update)
    model.iterations = 700
    oracle = O.OracleTransformer(**hp).load(model.get_weights_dict()); oracle.dropout_masks = masks
    lg_eval, _ = model([inp, tgt[:, :-1]], training=False)
    ref_eval, _ = oracle([torch.tensor(inp), torch.tensor(tgt[:, :-1])], training=False)
    assert _rel(lg_eval.cpu().numpy(), ref_eval.detach().numpy()) < 2e-2
    lg_train, _ = model([inp, tgt[:, :-1]], training=True)
    ref_train, _ = oracle([torch.tensor(inp), torch.tensor(tgt[:, :-1])], training=True)
    assert _rel(lg_train.cpu().numpy(), ref_train.detach().numpy()) < 2e-2
    assert _rel(lg_train.cpu().numpy(), lg_eval.cpu().numpy()) > 5e-2
