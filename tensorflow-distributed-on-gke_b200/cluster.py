"""Single-box launcher in place of the reference's Kubernetes bootstrap.

`TensorflowKubernetesCluster(worker_count)` (`cluster/cluster.py:12-79`) finds its peers through the Kubernetes API and a
heartbeat port, writes `TF_CONFIG`, builds `MultiWorkerMirroredStrategy()` (`:66`) and answers `chief()` (`:74-79`); when
`THIS_POD_NAME` is unset it runs in local mode with one worker (`:35-40`).  On one B200 box the peers are the processes
torchrun (or `launch` below) starts, one per GPU: rank and world size come from `RANK` / `WORLD_SIZE`, the rendezvous is
`torch.distributed` on 127.0.0.1, and the NCCL unique id travels through it when a model is attached to the strategy.

    cluster = B200Cluster(worker_count)          # TensorflowKubernetesCluster(WORKER_COUNT, verbose=True)
    with cluster.execution_strategy().scope():   # unchanged from __main__.py:37
        ...
    if cluster.chief(): save_snapshot()          # __main__.py:139-147,165-169
"""
import os
import socket

from . import _lib
from .training import B200Strategy


def _free_port():
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class B200Cluster:
    def __init__(self, worker_count, backend=None, master_addr="127.0.0.1", master_port=None, verbose=False):
        """worker_count: how many workers the job has (cluster/cluster.py:23-24).  Without RANK in the environment the
        process runs in local mode -- one worker, no rendezvous -- like the reference without THIS_POD_NAME (:35-40)."""
        self._verbose = verbose
        self._owns_group = False
        if "RANK" not in os.environ:
            print("Environment variable RANK not set, running in local mode.")
            self._rank, self._world = 0, 1
        else:
            self._rank = int(os.environ["RANK"])
            self._world = int(os.environ.get("WORLD_SIZE", "1"))
            if self._world != int(worker_count):
                raise _lib.B200XFError(f"worker_count={worker_count} but the launcher started WORLD_SIZE={self._world} processes")
            if not 0 <= self._rank < self._world:
                raise _lib.B200XFError(f"RANK={self._rank} outside WORLD_SIZE={self._world}")
            if self._world > 1:
                import torch
                import torch.distributed as dist
                if not dist.is_initialized():
                    os.environ.setdefault("MASTER_ADDR", master_addr)
                    if master_port is not None:
                        os.environ.setdefault("MASTER_PORT", str(master_port))
                    if "MASTER_PORT" not in os.environ:
                        raise _lib.B200XFError("MASTER_PORT is not set (torchrun and cluster.launch set it)")
                    use_cuda = torch.cuda.is_available()
                    backend = backend or ("nccl" if use_cuda else "gloo")
                    if use_cuda:
                        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", self._rank)) % max(torch.cuda.device_count(), 1))
                    if verbose:
                        print(f"worker {self._rank}: waiting for {self._world - 1} peers at "
                              f"{os.environ['MASTER_ADDR']}:{os.environ['MASTER_PORT']} ({backend})")
                    dist.init_process_group(backend, rank=self._rank, world_size=self._world)
                    self._owns_group = True
        self._execution_strategy = B200Strategy(self._rank, self._world)

    def execution_strategy(self):
        """The stand-in for tf.distribute.MultiWorkerMirroredStrategy (cluster/cluster.py:68-72)."""
        return self._execution_strategy

    def chief(self):
        """Is this process the chief -- worker 0, the one that writes snapshots (cluster/cluster.py:74-79)?"""
        return self._rank == 0

    @property
    def worker_index(self):
        return self._rank

    @property
    def worker_count(self):
        return self._world

    def shutdown(self):
        if self._owns_group:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.destroy_process_group()
            self._owns_group = False


def _worker(rank, world, port, fn, args):
    os.environ["RANK"] = str(rank)
    os.environ["LOCAL_RANK"] = str(rank)
    os.environ["WORLD_SIZE"] = str(world)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    fn(*args)


def launch(worker_count, fn, args=(), port=None):
    """Starts `worker_count` processes on this box (one per GPU), each with RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* set as
    torchrun would, runs fn(*args) in every one and returns when all have exited; raises if any of them failed.
    `fn` must be importable (defined at module level).  worker_count == 1 runs fn in this process in local mode."""
    if worker_count < 1:
        raise ValueError("worker_count must be at least 1")
    if worker_count == 1:
        return fn(*args)
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(worker_count, port or _free_port(), fn, tuple(args)), nprocs=worker_count, join=True)
