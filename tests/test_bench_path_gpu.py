"""Step-level parity of the kernels the BENCHMARKED configurations run (head size 64): the tcgen05 attention forward /
backward (single tile at S <= 128, multi-tile backward beyond), the cta_group::2 GEMM of the 32 000-wide vocabulary
projection, the d >= 512 LayerNorm instantiations and the Q/K/V / weight prefetch ahead of the dependency wait — checked
against the CPU fp32 oracle as whole train steps (logits, loss, whole gradient, weights after several Adam steps), with the
same rules and tolerances as tests/test_train_step_gpu.py (which covers the reference's tutorial geometry, head size 16).
Plus the orderings the captured CUDA graph of a step must survive (the advisor's cases): a forward_backward call between
replayed train steps, and communicator attach / detach."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import b200xf
from oracle import transformer_oracle as O

# Transformer-base layer geometry (BASELINE.json configs[2]) at reduced depth and batch: d 512, 8 heads of 64, dff 2048,
# vocabulary 32 000; 4 x 128 = 512 tokens give the vocabulary GEMMs their three waves of 256-wide tiles (CTA pairs).
BASE2 = dict(layers_count=2, embedding_size=512, heads_count=8, feed_forward_hidden_size=2048, input_vocabulary_size=32000,
             target_vocabulary_size=32000, maximum_input_length=1000, maximum_target_length=1000, droput_rate=0.0)
# Transformer-big layer geometry (configs[3]): d 1024, 16 heads of 64, dff 4096, S = T = 256 (two key tiles: multi-tile
# attention), smaller vocabulary to keep the CPU oracle quick.
BIG1 = dict(layers_count=1, embedding_size=1024, heads_count=16, feed_forward_hidden_size=4096, input_vocabulary_size=4000,
            target_vocabulary_size=4000, maximum_input_length=1000, maximum_target_length=1000, droput_rate=0.0)
CASES = [pytest.param(BASE2, 4, 128, 128, id="base-geometry"), pytest.param(BIG1, 2, 256, 256, id="big-geometry")]


def _rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


def _pair(hp, seed=0, emulate_bf16=False):
    model = b200xf.Transformer(**hp, seed=seed)
    return model, O.OracleTransformer(**hp, emulate_bf16=emulate_bf16).load(model.get_weights_dict())


@pytest.mark.parametrize("hp,B,S,T", CASES)
def test_logits_loss_and_gradients_parity_head_size_64(hp, B, S, T):
    torch.set_num_threads(os.cpu_count() or 1)
    model, oracle = _pair(hp)
    _, oracle16 = _pair(hp, emulate_bf16=True)
    vin, vt = hp["input_vocabulary_size"], hp["target_vocabulary_size"]
    inp, tgt = b200xf.synthetic_batch(B, S, T + 1, vin, vt, seed=4321, ragged=True)
    logits, _ = model([inp, tgt[:, :-1]], training=False)
    ref, _ = oracle([torch.tensor(inp), torch.tensor(tgt[:, :-1])])
    assert _rel(logits.cpu().numpy(), ref.detach().numpy()) < 2e-2                       # logits rel-L2 <= 2e-2
    out = model.forward_backward_device(inp, tgt, workers_count=2).cpu().numpy()
    loss_ref, logits_ref, grads_ref = O.train_step(oracle, inp, tgt, workers_count=2)
    _, _, grads16 = O.train_step(oracle16, inp, tgt, workers_count=2)
    assert abs(out[0] - loss_ref) < 1e-2 / 2, (out[0], loss_ref)                         # loss |delta| <= 1e-2 (/ workers)
    num, den = O.accuracy_counts(torch.tensor(tgt[:, 1:]), logits_ref)
    assert int(out[2]) == den and abs(int(out[1]) - num) <= max(2, den // 200)
    grads = model.get_slot_dict("grad")
    names = [n for n in grads_ref if not n.endswith("key_generator_weights/bias")]      # exactly 0 in real arithmetic
    cat = lambda g: np.concatenate([g[n].ravel() for n in names])
    e, e16 = _rel(cat(grads), cat(grads_ref)), _rel(cat(grads16), cat(grads_ref))
    assert e < 5e-2 and e < 1.3 * e16 + 5e-3, (e, e16)                                   # whole-gradient rel-L2; vs bf16 rounding alone
    worst = max((_rel(grads[n], grads_ref[n]), n) for n in names if np.linalg.norm(grads_ref[n]) > 1e-12)
    assert worst[0] < 0.2, worst                                                         # per-tensor rel-L2


@pytest.mark.parametrize("hp,B,S,T", CASES)
def test_train_steps_parity_head_size_64(hp, B, S, T):
    """5 Adam steps of one shape (so steps 3+ are CUDA-graph replays) against the oracle: loss at every step, delta-w."""
    torch.set_num_threads(os.cpu_count() or 1)
    model, oracle = _pair(hp)
    oracle16 = O.OracleTransformer(**hp, emulate_bf16=True).load(model.get_weights_dict())
    vin, vt = hp["input_vocabulary_size"], hp["target_vocabulary_size"]
    w0 = model.get_weights_dict()
    model.iterations = 2000; oracle.iterations = 2000; oracle16.iterations = 2000
    for step in range(5):
        inp, tgt = b200xf.synthetic_batch(B, S, T + 1, vin, vt, seed=900 + step, ragged=True)
        res = model.train_step_device(inp, tgt, 1).cpu().numpy()
        ref_loss, _, gref = O.train_step(oracle, inp, tgt, 1)
        O.train_step(oracle16, inp, tgt, 1)
        assert abs(res[0] - ref_loss) < 2e-2, (step, res[0], ref_loss)                   # loss curve within 2e-2 at every step
    assert model.iterations == 2005

    def delta_w_error(other):
        num = den = agree = tot = 0.0
        for name, p in oracle.params.items():
            ref = p.detach().numpy(); d_ref = ref - w0[name]; d_got = other[name] - w0[name]
            sel = np.abs(gref[name]) > 1e-7
            if sel.any():
                num += float(np.sum((d_got[sel] - d_ref[sel]) ** 2)); den += float(np.sum(d_ref[sel] ** 2))
                wgt = np.abs(gref[name][sel])
                agree += float(np.sum(wgt * (np.sign(d_got[sel]) == np.sign(d_ref[sel])))); tot += float(np.sum(wgt))
        return np.sqrt(num / den), agree / tot
    err, sign = delta_w_error(model.get_weights_dict())
    err16, sign16 = delta_w_error({k: p.detach().numpy() for k, p in oracle16.params.items()})
    assert err < 0.15 and err < 1.3 * err16 + 0.01, (err, err16)                         # same rule as test_train_steps_parity
    assert sign > min(0.99, sign16 - 0.005), (sign, sign16)


def test_forward_backward_between_replayed_train_steps():
    """train x4 (the 3rd and 4th replay the captured graph), forward_backward (leaves gradients behind), train: the replayed
    step must not add onto the stale gradients.  Twin model without the forward_backward call; and the oracle."""
    hp = dict(BASE2, layers_count=1, input_vocabulary_size=2000, target_vocabulary_size=2000)
    a = b200xf.Transformer(**hp, seed=5); b = b200xf.Transformer(**hp, seed=5)
    oracle = O.OracleTransformer(**hp).load(a.get_weights_dict())
    for m in (a, b, oracle):
        m.iterations = 1500
    batches = [b200xf.synthetic_batch(4, 64, 65, 2000, 2000, seed=50 + i, ragged=True) for i in range(6)]
    for i in range(4):
        a.train_step_device(*batches[i], 1); b.train_step_device(*batches[i], 1); O.train_step(oracle, *batches[i], 1)
    a.forward_backward_device(*batches[5], 1)
    assert float(a._flat("grad").abs().sum()) > 0
    ra = a.train_step_device(*batches[4], 1).cpu().numpy(); rb = b.train_step_device(*batches[4], 1).cpu().numpy()
    ref_loss, _, _ = O.train_step(oracle, *batches[4], 1)
    np.testing.assert_allclose(ra[:3], rb[:3], rtol=1e-5)
    assert abs(ra[0] - ref_loss) < 2e-2
    wa, wb = a.get_weights_dict(), b.get_weights_dict()
    assert max(_rel(wa[k], wb[k]) for k in wa) < 1e-5      # identical up to the order of the split-K reduce-adds
    assert float(a._flat("grad").abs().sum()) == 0.0       # every step leaves the gradients cleared


def _one_rank_comm(model):
    buf = C.create_string_buffer(128)
    b200xf._lib.check(b200xf._lib.load().b200xf_comm_unique_id(buf))
    model.comm_init(bytes(buf.raw), 0, 1)


def test_graph_is_dropped_when_a_communicator_is_attached_or_detached():
    hp = dict(BASE2, layers_count=1, input_vocabulary_size=2000, target_vocabulary_size=2000)
    a = b200xf.Transformer(**hp, seed=6); b = b200xf.Transformer(**hp, seed=6)
    a.iterations = 800; b.iterations = 800
    batches = [b200xf.synthetic_batch(4, 64, 65, 2000, 2000, seed=80 + i, ragged=True) for i in range(9)]
    lib = b200xf._lib.load()
    for i in range(9):
        if i == 3:
            _one_rank_comm(a)                               # the graph captured at step 2 has no collectives in it
        if i == 6:
            b200xf._lib.check(lib.b200xf_comm_destroy(a._ctx))  # the graph captured at step 5 holds this communicator's
        ra = a.train_step_device(*batches[i], 1).cpu().numpy(); rb = b.train_step_device(*batches[i], 1).cpu().numpy()
        np.testing.assert_allclose(ra, rb, rtol=1e-5)
    wa, wb = a.get_weights_dict(), b.get_weights_dict()
    assert max(_rel(wa[k], wb[k]) for k in wa) < 1e-5


UPDATE_MODES = [dict(B200XF_SHARDED_UPDATE="0"), dict(B200XF_SHARDED_UPDATE="2", B200XF_PIPELINED_UPDATE="0"),
                dict(B200XF_SHARDED_UPDATE="2", B200XF_PIPELINED_UPDATE="1"), dict(B200XF_SHARDED_UPDATE="2", B200XF_PIPELINED_UPDATE="2")]


def test_sharded_pipelined_update_chain_on_one_rank(monkeypatch):
    """The update path the multi-GPU runs take — reduce-scatter, Adam on the shard, all-gather, bf16 refresh, with the
    trailing (or all) buckets pipelined over two streams and alternating event slots — driven on ONE GPU through a
    one-rank communicator (B200XF_SHARDED_UPDATE=2), several 1 MB buckets, 6 steps of one shape so that the captured graph
    replays: every variant must produce the weights of the plain single-GPU step and of the replicated all-reduce update."""
    hp = dict(layers_count=3, embedding_size=128, heads_count=2, feed_forward_hidden_size=512, input_vocabulary_size=5000,
              target_vocabulary_size=5000, maximum_input_length=200, maximum_target_length=200, droput_rate=0.0)
    batches = [b200xf.synthetic_batch(8, 48, 49, 5000, 5000, seed=300 + i, ragged=True) for i in range(6)]

    def run(env, comm):
        for k in ("B200XF_SHARDED_UPDATE", "B200XF_PIPELINED_UPDATE", "B200XF_BUCKET_MB"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        monkeypatch.setenv("B200XF_BUCKET_MB", "1")
        m = b200xf.Transformer(**hp, seed=9)
        m.trainable_variables
        if comm:
            _one_rank_comm(m)
        m.iterations = 1200
        res = [m.train_step_device(*bt, 1).cpu().numpy() for bt in batches]
        w = m.get_weights_dict(); mm = m.get_slot_dict("adam_m")
        m.close()
        return np.stack(res), w, mm
    ref_res, ref_w, ref_m = run({}, comm=False)
    for env in UPDATE_MODES:
        res, w, mm = run(env, comm=True)
        np.testing.assert_allclose(res, ref_res, rtol=2e-5, err_msg=str(env))
        assert max(_rel(w[k], ref_w[k]) for k in w) < 1e-5, env        # up to the order of the split-K reduce-adds
        assert max(_rel(mm[k], ref_m[k]) for k in mm if np.linalg.norm(ref_m[k]) > 0) < 1e-4, env
