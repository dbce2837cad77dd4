"""world_size-2 `gloo` tests (CPU) of the host-side data-parallel logic: the strategy object's dataset sharding and
SUM reduce, and — with the oracle as the arithmetic — the reference's gradient semantics the NCCL path must
reproduce: each replica normalises by its OWN non-pad count and divides by W, gradients are SUMmed
(transformer_model.py:16-17, __main__.py:113-114,131-132)."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

HP = dict(layers_count=1, embedding_size=32, heads_count=2, feed_forward_hidden_size=64, input_vocabulary_size=50,
          target_vocabulary_size=40, maximum_input_length=32, maximum_target_length=32, droput_rate=0.0)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import b200xf
    from oracle import transformer_oracle as O
    strategy = b200xf.B200Strategy()
    assert strategy.num_replicas_in_sync == world and strategy.rank == rank
    # global batches of 2*world sentences; every replica keeps its own contiguous slice (AutoShardPolicy.DATA)
    glob = [b200xf.synthetic_batch(2 * world, 8, 9, 50, 40, seed=10 + i, ragged=True) for i in range(2)]
    shards = list(strategy.experimental_distribute_dataset(glob))
    assert all(np.array_equal(s[0], g[0][2 * rank:2 * rank + 2]) for s, g in zip(shards, glob))
    # per-replica oracle step with a gloo SUM all-reduce of the gradients
    model = O.OracleTransformer(**HP).load(O.init_parameters(1, 32, 64, 50, 40, seed=0))
    model.iterations = 100

    def allreduce(grads):
        for k in grads:
            t = torch.from_numpy(grads[k]); dist.all_reduce(t); grads[k] = t.numpy()
        return grads
    loss, _, grads = O.train_step(model, shards[0][0], shards[0][1], workers_count=world, allreduce=allreduce)
    total = strategy.reduce(b200xf.ReduceOp.SUM, torch.tensor([loss]), axis=None)
    if rank == 0:
        np.savez(out, loss_sum=total.numpy(), **{k.replace("/", "|"): v for k, v in grads.items()},
                 **{"W|" + k.replace("/", "|"): p.detach().numpy() for k, p in model.params.items()})
    dist.barrier()
    dist.destroy_process_group()


def test_two_replicas_match_mean_of_local_means(tmp_path):
    world, port, out = 2, _free_port(), str(tmp_path / "r0.npz")
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    z = np.load(out)
    import b200xf
    from oracle import transformer_oracle as O
    glob = b200xf.synthetic_batch(2 * world, 8, 9, 50, 40, seed=10, ragged=True)
    # single-process restatement: SUM over replicas of grad(local_loss / W)
    want = None; loss_sum = 0.0
    for r in range(world):
        m = O.OracleTransformer(**HP).load(O.init_parameters(1, 32, 64, 50, 40, seed=0)); m.iterations = 100
        loss, _, g = O.train_step(m, glob[0][2 * r:2 * r + 2], glob[1][2 * r:2 * r + 2], workers_count=world)
        loss_sum += loss
        want = g if want is None else {k: want[k] + g[k] for k in g}
    np.testing.assert_allclose(z["loss_sum"][0], loss_sum, rtol=1e-5)
    for k, v in want.items():
        np.testing.assert_allclose(z[k.replace("/", "|")], v, rtol=2e-4, atol=1e-8)
    # and it is NOT the global token mean unless the shards hold equally many tokens
    m = O.OracleTransformer(**HP).load(O.init_parameters(1, 32, 64, 50, 40, seed=0))
    _, _, gglob = O.train_step(m, glob[0], glob[1], workers_count=1)
    k = "final_layer/kernel"
    assert np.linalg.norm(gglob[k] - want[k]) / np.linalg.norm(want[k]) > 1e-3


def test_strategy_single_replica_is_identity():
    import b200xf
    s = b200xf.B200Strategy(0, 1)
    with s.scope():
        assert s.run(lambda a, b: a + b, args=(1, 2)) == 3
    t = torch.tensor([1.5])
    assert s.reduce(b200xf.ReduceOp.SUM, t, axis=None) is t
    data = [(np.arange(4), np.arange(4))]
    assert list(s.experimental_distribute_dataset(data))[0][0].tolist() == [0, 1, 2, 3]
