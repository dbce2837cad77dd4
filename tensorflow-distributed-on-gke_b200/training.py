"""Host-side mirror of the training pieces of the reference's `__main__.py` (schedule, optimizer,
train_step / distributed_train_step / test_step) and of the strategy object it uses
(cluster/cluster.py:66-79), over libb200xf.so and NCCL.
"""
from __future__ import annotations

import contextlib
import os

import numpy as np
import torch

from . import _lib
from ._lib import check


class TransformerLearningRateSchedule:
    """__main__.py:56-67 — lr = rsqrt(d) * min(step * warmup^-1.5, rsqrt(step)), fp32.  The library evaluates the same
    expression on the device inside the fused Adam; this host copy exists for logging and tests."""

    def __init__(self, embedding_size, warmup_steps=4000):
        self.embedding_size = np.float32(embedding_size)
        self.warmup_steps = warmup_steps

    def __call__(self, step):
        step = np.float32(step)
        with np.errstate(divide="ignore"):
            rising = step * np.float32(float(self.warmup_steps) ** -1.5)
            falling = np.float32(1.0) / np.sqrt(step, dtype=np.float32)
            return np.float32(1.0) / np.sqrt(self.embedding_size, dtype=np.float32) * np.minimum(rising, falling)


CustomSchedule = TransformerLearningRateSchedule  # TF-tutorial name (BASELINE.json)


class Adam:
    """tf.keras.optimizers.Adam(learning_rate, beta_1, beta_2, epsilon) as used at __main__.py:72-73.

    Only a TransformerLearningRateSchedule learning rate is supported (that is the path being replaced); the
    update itself is the library's fused TF-form Adam over the model's flat buffers."""

    def __init__(self, learning_rate, beta_1=0.9, beta_2=0.999, epsilon=1e-7):
        if not isinstance(learning_rate, TransformerLearningRateSchedule):
            raise TypeError("Adam here requires a TransformerLearningRateSchedule learning rate")
        self.learning_rate, self.beta_1, self.beta_2, self.epsilon = learning_rate, beta_1, beta_2, epsilon
        self._model = None

    def bind(self, model):
        ok = (model._opt["adam_beta_1"] == float(self.beta_1) and model._opt["adam_beta_2"] == float(self.beta_2)
              and model._opt["adam_epsilon"] == float(self.epsilon) and model._opt["warmup_steps"] == int(self.learning_rate.warmup_steps)
              and model.hparams["embedding_size"] == int(self.learning_rate.embedding_size))
        if not ok:
            if model._ctx is not None:
                raise _lib.B200XFError("optimizer constants differ from the ones the model was built with")
            model._opt.update(adam_beta_1=float(self.beta_1), adam_beta_2=float(self.beta_2), adam_epsilon=float(self.epsilon),
                              warmup_steps=int(self.learning_rate.warmup_steps))
        self._model = model
        return self

    @property
    def iterations(self):
        return self._model.iterations if self._model is not None else 0

    def apply_gradients(self, grads_and_vars):
        """__main__.py:113-114.  Accepts the (gradient, variable) pairs of `Transformer` variables whose gradients were
        produced by `forward_backward_device`; applies the fused Adam over the whole flat buffer and bumps iterations."""
        pairs = list(grads_and_vars)
        if not pairs:
            return
        model = pairs[0][1]._model
        self.bind(model)
        from .kernels import adam_step
        it = torch.tensor([model.iterations], device=model.device, dtype=torch.int64)
        adam_step(model._flat("param"), model._flat("grad"), model._flat("adam_m"), model._flat("adam_v"), it,
                  model.hparams["embedding_size"], self.learning_rate.warmup_steps, self.beta_1, self.beta_2, self.epsilon)
        model._dirty = True
        model._sync_weights()
        model.iterations = model.iterations + 1


class ReduceOp:
    SUM = "SUM"
    MEAN = "MEAN"


class B200Strategy:
    """Stand-in for tf.distribute.MultiWorkerMirroredStrategy() (cluster/cluster.py:66): one process per GPU,
    NCCL over NVLink inside the library.  Rendezvous uses torch.distributed (env:// by default) only to broadcast
    the NCCL unique id."""

    def __init__(self, rank=None, world_size=None):
        self.rank = int(os.environ.get("RANK", 0)) if rank is None else rank
        self.world_size = int(os.environ.get("WORLD_SIZE", 1)) if world_size is None else world_size
        self._models = []

    @property
    def num_replicas_in_sync(self):
        return self.world_size

    @contextlib.contextmanager
    def scope(self):
        yield self

    def attach(self, model):
        """Give `model` a communicator (no-op for a single replica)."""
        if self.world_size > 1 and model._comm is None:
            import torch.distributed as dist
            if not dist.is_initialized():
                raise _lib.B200XFError("torch.distributed must be initialised before attaching a multi-replica strategy")
            uid = [None]
            if self.rank == 0:
                import ctypes as C
                buf = C.create_string_buffer(128)
                check(_lib.load().b200xf_comm_unique_id(buf))
                uid[0] = bytes(buf.raw)
            dist.broadcast_object_list(uid, src=0)
            model.comm_init(uid[0], self.rank, self.world_size)
            # One fused kernel per bucket (reduce the rank's gradient shard over all replicas + Adam + write the new weights to every
            # replica) over symmetric memory.  Through the NVSwitch (multimem) it moves ~(1 + 1/W) bucket sizes per link direction,
            # point to point (W-1)/W (profiles/r02_nvls_update.md).
            # B200XF_NVLS = auto | 1 (multimem) | p2p (point to point) | 0 (NCCL reduce-scatter / all-gather).
            mode = os.environ.get("B200XF_NVLS", "auto")
            if mode == "auto":
                mode = "1" if self.world_size >= 4 else "0"
            if mode in ("1", "p2p"):
                self._attach_nvls(model, point_to_point=(mode == "p2p"))
        self._models.append(model)
        return model

    def _attach_nvls(self, model, point_to_point=False):
        """Move the model's flat weights and gradients into symmetric memory with a multicast (NVLS) mapping and switch the
        gradient exchange to the fused multimem kernel (reduce-scatter + Adam + all-gather in one launch).  Falls back to the
        NCCL path when the box offers no multicast mapping (B200XF_NVLS=0 forces that)."""
        import ctypes as C
        import torch.distributed as dist
        if self.world_size > 8:
            return False
        try:
            import torch.distributed._symmetric_memory as symm_mem
            lib = _lib.load()
            total = int(lib.b200xf_param_total(model._ctx))
            dev = model.device
            group = dist.group.WORLD.group_name
            bufs = [symm_mem.empty(total, dtype=torch.float32, device=dev) for _ in range(2)]
            pad = symm_mem.empty(256, dtype=torch.int32, device=dev)
            handles = [symm_mem.rendezvous(t, group) for t in bufs + [pad]]
            ok = point_to_point or all(int(h.multicast_ptr) != 0 for h in handles[:2])
        except Exception as e:  # no symmetric memory on this build / box
            ok = False; why = repr(e)
        flag = torch.tensor([1 if ok else 0], device=model.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)           # all replicas take the same path
        if int(flag.item()) == 0:
            if self.rank == 0 and os.environ.get("B200XF_NVLS") == "1":
                print(f"[b200xf] NVLS update unavailable, using NCCL: {locals().get('why', 'no multicast mapping')}", flush=True)
            return False
        for t in bufs:
            t.zero_()
        pad.zero_()
        check(lib.b200xf_use_external_state(model._ctx, C.c_void_p(bufs[0].data_ptr()), C.c_void_p(bufs[1].data_ptr())))
        model._bind(model._ctx)                                  # the variable views now wrap the symmetric buffers
        torch.cuda.synchronize(model.device)
        dist.barrier()                                           # every replica's buffers and flag pad are initialised
        arr = lambda h: (C.c_void_p * self.world_size)(*[int(p) for p in h.buffer_ptrs])
        check(lib.b200xf_nvls_attach(model._ctx, C.c_void_p(int(handles[1].multicast_ptr) or None), C.c_void_p(int(handles[0].multicast_ptr) or None),
                                     arr(handles[2]), arr(handles[1]), arr(handles[0]), int(point_to_point), self.rank, self.world_size))
        model._symmetric = (bufs, pad, handles)                  # keep the allocations alive as long as the model
        return True

    def run(self, fn, args=(), kwargs=None):
        return fn(*args, **(kwargs or {}))

    def reduce(self, reduce_op, value, axis=None):
        """strategy.reduce(SUM, per_replica_losses, axis=None), __main__.py:131-132.  `value` may be the 6-vector a
        train step returns (already reduced inside the step: entries 3..5) or a plain tensor."""
        if isinstance(value, StepResult):
            return value.reduced_loss
        if self.world_size == 1:
            return value
        import torch.distributed as dist
        t = value.clone()
        dist.all_reduce(t)
        return t / self.world_size if reduce_op == ReduceOp.MEAN else t

    def experimental_distribute_dataset(self, dataset):
        """AutoShardPolicy.DATA (english_portugese_dataset.py:27-31): every replica walks the whole dataset of global
        batches and keeps its own contiguous slice of each."""
        rank, world = self.rank, self.world_size

        from .batching import replica_rows

        def gen():
            for inp, tgt in dataset:
                first, n = replica_rows(len(inp), rank, world)  # uneven batches: sizes differ by at most one, none dropped
                yield inp[first:first + n], tgt[first:first + n]
        return gen()


class StepResult:
    """What train_step returns: behaves like the local loss scalar, carries the cross-replica sums."""

    def __init__(self, vec):
        self.vec = vec

    @property
    def loss(self): return self.vec[0]
    @property
    def reduced_loss(self): return self.vec[3]
    @property
    def accuracy(self): return self.vec[1] / self.vec[2]

    def __float__(self): return float(self.vec[0])


class Mean:
    """tf.keras.metrics.Mean (__main__.py:91-93)."""

    def __init__(self, name=None):
        self.name = name; self.reset_states()

    def reset_states(self):
        self.total = 0.0; self.count = 0

    def update_state(self, v):
        self.total += float(v); self.count += 1

    def result(self):
        return self.total / max(self.count, 1)


def make_train_step(transformer, optimizer, workers_count, train_accuracy=None):
    """Builds the reference's `train_step(input, target)` (__main__.py:105-117) as one fused library call."""
    optimizer.bind(transformer)

    def train_step(input, target):
        res = StepResult(transformer.train_step_device(input, target, workers_count))
        if train_accuracy is not None:
            train_accuracy.update_state(res.accuracy)
        return res
    return train_step


def make_distributed_train_step(strategy, train_step):
    """`distributed_train_step` (__main__.py:127-132): strategy.run + strategy.reduce(SUM)."""
    def distributed_train_step(input, target):
        per_replica = strategy.run(train_step, args=(input, target))
        return strategy.reduce(ReduceOp.SUM, per_replica, axis=None)
    return distributed_train_step


def make_test_step(transformer, test_loss=None, test_accuracy=None):
    """`test_step` (__main__.py:119-125): forward, loss with workers_count=1, accuracy; no update."""
    def test_step(input, target):
        res = StepResult(transformer.eval_step_device(input, target))
        if test_loss is not None: test_loss.update_state(res.loss)
        if test_accuracy is not None: test_accuracy.update_state(res.accuracy)
        return res
    return test_step


def make_distributed_test_step(strategy, test_step):
    """`distributed_test_step` (__main__.py:134-137): strategy.run(test_step) on this replica's shard; like the reference it
    returns nothing — the metrics objects given to make_test_step carry the result."""
    def distributed_test_step(input, target):
        strategy.run(test_step, args=(input, target))
    return distributed_test_step
