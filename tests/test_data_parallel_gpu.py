"""Data-parallel path on the GPU: bucketed NCCL all-reduce with the Adam update chained behind each bucket.
The 1-rank communicator case runs on any box; the 2-rank case needs two GPUs (skipped otherwise) and checks the
reference's semantics (SUM of gradients of local_loss / W) against the CPU oracle."""
import ctypes as C
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import b200xf
from oracle import transformer_oracle as O

HP = dict(layers_count=2, embedding_size=64, heads_count=4, feed_forward_hidden_size=128, input_vocabulary_size=203,
          target_vocabulary_size=157, maximum_input_length=64, maximum_target_length=64, droput_rate=0.0)


def _rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


def test_single_rank_communicator_matches_plain_step():
    lib = b200xf._lib.load()
    a = b200xf.Transformer(**HP, seed=1); b = b200xf.Transformer(**HP, seed=1)
    a.trainable_variables; b.trainable_variables
    buf = C.create_string_buffer(128)
    b200xf._lib.check(lib.b200xf_comm_unique_id(buf))
    b.comm_init(bytes(buf.raw), 0, 1)
    a.iterations = 300; b.iterations = 300
    for step in range(3):
        inp, tgt = b200xf.synthetic_batch(4, 20, 18, 203, 157, seed=40 + step, ragged=True)
        ra = a.train_step_device(inp, tgt, 1).cpu().numpy(); rb = b.train_step_device(inp, tgt, 1).cpu().numpy()
        np.testing.assert_allclose(ra, rb, rtol=1e-5)
        np.testing.assert_allclose(rb[3:], rb[:3], rtol=1e-6)  # one replica: the reduced metrics equal the local ones
    wa, wb = a.get_weights_dict(), b.get_weights_dict()
    assert max(_rel(wa[k], wb[k]) for k in wa) < 1e-5  # identical up to atomic split-K ordering
    assert a.iterations == b.iterations == 303


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


# several 1 MB buckets (final layer, decoder layers, embeddings + K|V, encoder layers): the RS(k+1) / AG(k) interleave of the
# pipelined sharded update and its alternating event slots run; 6 steps of one shape, so steps 3+ replay the captured graph
HP2 = dict(layers_count=3, embedding_size=128, heads_count=2, feed_forward_hidden_size=512, input_vocabulary_size=5000,
           target_vocabulary_size=5000, maximum_input_length=200, maximum_target_length=200, droput_rate=0.0)
STEPS2, B2, S2, T2 = 6, 8, 48, 48


def _rank_main(rank, world, port, out_dir, env):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    os.environ.update(env)
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    model = b200xf.Transformer(**HP2, seed=2, device=f"cuda:{rank}")
    strategy = b200xf.B200Strategy(rank, world)
    with strategy.scope():
        model.trainable_variables
        strategy.attach(model)
    model.iterations = 500
    res = []
    for step in range(STEPS2):
        glob = b200xf.synthetic_batch(B2 * world, S2, T2 + 1, 5000, 5000, seed=70 + step, ragged=True)
        inp, tgt = next(iter(strategy.experimental_distribute_dataset([glob])))
        res.append(model.train_step_device(inp, tgt, world).cpu().numpy())
    w = model.get_weights_dict()
    sharded = os.environ.get("B200XF_SHARDED_UPDATE", "1") != "0"
    raised = False
    try:
        model.get_slot_dict("adam_m")
    except b200xf.B200XFError:
        raised = True
    assert raised == sharded, "sharded Adam slots must not be readable before gather_optimizer_state()"
    model.gather_optimizer_state()
    m = model.get_slot_dict("adam_m"); v = model.get_slot_dict("adam_v")
    assert b200xf._lib.load().b200xf_nvls_error(model._ctx) == 0
    nvls = model._symmetric is not None
    # dropout: every replica draws its own masks (the rank is part of the key)
    hp_drop = dict(HP2, droput_rate=0.1)
    dm = b200xf.Transformer(**hp_drop, seed=2, device=f"cuda:{rank}")
    dm.trainable_variables
    strategy.attach(dm)
    mask = dm.dropout_mask(2, 64).cpu().numpy()
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), res=np.stack(res), mask=mask, nvls=np.array([int(nvls)]),
             **{"w|" + k.replace("/", "|"): a for k, a in w.items()}, **{"m|" + k.replace("/", "|"): a for k, a in m.items()},
             **{"v|" + k.replace("/", "|"): a for k, a in v.items()})
    dist.barrier()
    dm.close(); model.close()
    dist.destroy_process_group()


def _oracle_run(world):
    w0 = b200xf.Transformer(**HP2, seed=2).get_weights_dict()
    oracle = O.OracleTransformer(**HP2).load(w0); oracle.iterations = 500
    per_step = []
    for step in range(STEPS2):
        glob = b200xf.synthetic_batch(B2 * world, S2, T2 + 1, 5000, 5000, seed=70 + step, ragged=True)
        shards = [(glob[0][B2 * r:B2 * (r + 1)], glob[1][B2 * r:B2 * (r + 1)]) for r in range(world)]
        losses, grads = [], None
        for r in range(world):  # W replicas, each loss / W on its own shard (transformer_model.py:16-17), gradients SUMmed
            probe = O.OracleTransformer(**HP2).load({k: p.detach().numpy() for k, p in oracle.params.items()})
            probe.iterations = oracle.iterations
            l, _, g = O.train_step(probe, *shards[r], workers_count=world)
            losses.append(l); grads = g if grads is None else {k: grads[k] + g[k] for k in g}
        O.train_step(oracle, *shards[0], workers_count=world, allreduce=lambda _g: grads)
        per_step.append(losses)
    return w0, oracle, grads, per_step


MODES = [pytest.param({"B200XF_NVLS": "1"}, id="nvls-fused-update"), pytest.param({"B200XF_NVLS": "p2p"}, id="p2p-fused-update"),
         pytest.param({"B200XF_NVLS": "0"}, id="nccl-sharded-pipelined-trailing"), pytest.param({"B200XF_NVLS": "0", "B200XF_PIPELINED_UPDATE": "1"}, id="nccl-pipelined-all"),
         pytest.param({"B200XF_NVLS": "0", "B200XF_PIPELINED_UPDATE": "0"}, id="nccl-sharded-chained"), pytest.param({"B200XF_SHARDED_UPDATE": "0"}, id="nccl-replicated-allreduce")]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("env", MODES)
def test_two_ranks_match_oracle_and_each_other(tmp_path, env):
    import torch.multiprocessing as mp
    world = 2
    env = dict(env, B200XF_BUCKET_MB="1")
    mp.spawn(_rank_main, args=(world, _free_port(), str(tmp_path), env), nprocs=world, join=True)
    z = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    if env.get("B200XF_NVLS") in ("1", "p2p"):
        assert int(z[0]["nvls"][0]) == 1, "the fused NVLS update was expected to be active on this box"
    names = [k for k in z[0].files if k[:2] in ("w|", "m|", "v|")]
    for k in names:  # mirrored variables and (after the gather) optimizer slots: identical on every replica
        assert np.array_equal(z[1][k], z[0][k]), k
    np.testing.assert_allclose(z[0]["res"][:, 3:], z[1]["res"][:, 3:], rtol=1e-6)  # reduced metrics agree on all ranks
    assert not np.array_equal(z[0]["mask"], z[1]["mask"])                          # independent dropout masks per replica
    w0, oracle, grads, per_step = _oracle_run(world)
    for step, losses in enumerate(per_step):
        assert abs(z[0]["res"][step, 3] - sum(losses)) < 2e-2          # strategy.reduce(SUM, per-replica losses), __main__.py:131-132
        assert abs(z[0]["res"][step, 0] - losses[0]) < 2e-2 and abs(z[1]["res"][step, 0] - losses[1]) < 2e-2
    num = den = 0.0
    for k, p in oracle.params.items():
        ref = p.detach().numpy() - w0[k]; got = z[0]["w|" + k.replace("/", "|")] - w0[k]
        sel = np.abs(grads[k]) > 1e-7
        num += float(np.sum((got[sel] - ref[sel]) ** 2)); den += float(np.sum(ref[sel] ** 2))
    assert np.sqrt(num / den) < 0.2, np.sqrt(num / den)  # delta-w rel-L2 (bf16 compute vs fp32 oracle, early-Adam steps)
    # Adam's first moment against the oracle's (fp32 sums of bf16-computed gradients): whole-state rel-L2
    mn = np.sqrt(sum(float(np.sum((z[0]["m|" + k.replace("/", "|")] - oracle.m[k]) ** 2)) for k in oracle.m))
    md = np.sqrt(sum(float(np.sum(oracle.m[k] ** 2)) for k in oracle.m))
    assert mn / md < 6e-2, mn / md


def _two_contexts_main(out):
    """One process, two GPUs, one context each (include/b200xf.h: "distinct contexts are independent"): every kernel that raises its
    dynamic shared memory limit must do so on BOTH devices (cudaFuncSetAttribute is per device)."""
    hp = dict(layers_count=1, embedding_size=512, heads_count=8, feed_forward_hidden_size=1024, input_vocabulary_size=3000,
              target_vocabulary_size=3000, maximum_input_length=300, maximum_target_length=300, droput_rate=0.0)
    res = []
    for dev in ("cuda:0", "cuda:1"):
        with torch.cuda.device(dev):
            m = b200xf.Transformer(**hp, seed=3, device=dev)
            m.iterations = 100
            for S in (128, 260):   # single-tile and multi-tile attention kernels, d = 512 LayerNorm, >48 KB GEMM configurations
                inp, tgt = b200xf.synthetic_batch(2, S, S + 1, 3000, 3000, seed=S, ragged=True)
                res.append(m.train_step_device(inp, tgt, 1).cpu().numpy())
            m.close()
    np.save(out, np.stack(res))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_contexts_on_two_gpus_in_one_process(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "two_ctx.npy")
    p = mp.get_context("spawn").Process(target=_two_contexts_main, args=(out,))
    p.start(); p.join(300)
    assert p.exitcode == 0
    r = np.load(out)
    assert np.isfinite(r).all()
    np.testing.assert_allclose(r[:2], r[2:], rtol=1e-5)   # the same model and batches give the same losses on either device


def _launched_worker(out_dir):
    """What a worker of the reference's __main__.py does, started by b200xf.launch instead of Kubernetes (cluster/cluster.py:12-79)."""
    cluster = b200xf.B200Cluster(2)
    strategy = cluster.execution_strategy()
    rank = cluster.worker_index
    with strategy.scope():
        model = b200xf.Transformer(**HP, seed=4, device=f"cuda:{rank}")
        optimizer = b200xf.Adam(b200xf.TransformerLearningRateSchedule(HP["embedding_size"]), beta_1=0.9, beta_2=0.98, epsilon=1e-9)
        b200xf.build_model(model)
        strategy.attach(model)
        step = b200xf.make_distributed_train_step(strategy, b200xf.make_train_step(model, optimizer, cluster.worker_count))
    model.iterations = 200
    batches = [b200xf.synthetic_batch(8, 20, 19, 203, 157, seed=90 + i, ragged=True) for i in range(3)]
    losses = [float(step(inp, tgt)) for inp, tgt in strategy.experimental_distribute_dataset(batches)]
    np.savez(os.path.join(out_dir, f"launched{rank}.npz"), losses=np.array(losses), chief=np.array([int(cluster.chief())]),
             w=np.concatenate([a.ravel() for a in model.get_weights_dict().values()]))
    model.close()
    cluster.shutdown()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_launcher_and_cluster_drive_two_gpus(tmp_path):
    b200xf.launch(2, _launched_worker, args=(str(tmp_path),))
    z = [np.load(tmp_path / f"launched{r}.npz") for r in range(2)]
    assert [int(f["chief"][0]) for f in z] == [1, 0]
    np.testing.assert_allclose(z[0]["losses"], z[1]["losses"], rtol=1e-6)   # the SUM over replicas, the same on both
    assert np.isfinite(z[0]["losses"]).all() and np.array_equal(z[0]["w"], z[1]["w"])
