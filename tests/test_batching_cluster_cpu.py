"""CPU tests of SURVEY §8f ranks 3 and 4: the host logic of on-device batch assembly (ragged -> padded, length bucketing,
per-replica split; english_portugese_dataset.py:36-47, __main__.py:48-53) and the single-box launcher that replaces the
Kubernetes bootstrap (cluster/cluster.py:12-79: local mode, chief(), worker_count check, rendezvous)."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import b200xf  # noqa: E402


def _ragged(n, lo, hi, vocab, seed):
    rng = np.random.default_rng(seed)
    return [list(rng.integers(1, vocab, size=int(rng.integers(lo, hi + 1)))) for _ in range(n)]


def test_ragged_to_flat_and_pad_to_longest_round_trip():
    sents = _ragged(7, 1, 9, 50, 0)
    flat, off = b200xf.ragged_to_flat(sents)
    assert off[0] == 0 and off[-1] == len(flat) == sum(len(s) for s in sents)
    for r, s in enumerate(sents):
        assert list(flat[off[r]:off[r + 1]]) == s
    dense = b200xf.pad_to_longest(sents)
    assert dense.shape == (7, max(len(s) for s in sents)) and dense.dtype == np.int64
    for r, s in enumerate(sents):
        assert list(dense[r, :len(s)]) == s and not dense[r, len(s):].any()  # PAD = 0 after the sentence, as .to_tensor()


def test_padded_length_bucketing_bounds_the_shapes():
    _, off = b200xf.ragged_to_flat(_ragged(5, 3, 37, 50, 1))
    longest = int(np.diff(off).max())
    assert b200xf.padded_length(off) == longest
    L16 = b200xf.padded_length(off, 16)
    assert L16 % 16 == 0 and longest <= L16 < longest + 16
    assert b200xf.padded_length(off, 16, maximum_length=longest) == longest  # never beyond the model's maximum
    with pytest.raises(ValueError):
        b200xf.padded_length(off, 16, maximum_length=longest - 1)
    with pytest.raises(ValueError):
        b200xf.padded_length(np.zeros(3, dtype=np.int64))
    # 200 random batches with lengths up to 100: at most ceil(100 / 16) distinct padded lengths
    seen = {b200xf.padded_length(b200xf.ragged_to_flat(_ragged(8, 1, 100, 50, 10 + i))[1], 16) for i in range(200)}
    assert len(seen) <= 7


def test_replica_rows_match_the_strategy_split():
    sents = _ragged(13, 2, 6, 50, 2)
    dense = b200xf.pad_to_longest(sents)
    for world in (1, 2, 3, 4):
        covered = []
        for rank in range(world):
            first, rows = b200xf.replica_rows(len(sents), rank, world)
            shard = next(iter(b200xf.B200Strategy(rank, world).experimental_distribute_dataset([(dense, dense)])))
            assert np.array_equal(shard[0], dense[first:first + rows])
            covered += list(range(first, first + rows))
        assert covered == list(range(len(sents)))  # contiguous, disjoint, every row kept (sizes differ by at most one)
    with pytest.raises(ValueError):
        b200xf.replica_rows(8, 2, 2)


def test_local_mode_without_rank(monkeypatch, capsys):
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    cluster = b200xf.B200Cluster(3)  # like the reference off-cluster: the worker count is ignored, one local worker
    assert "local mode" in capsys.readouterr().out
    assert cluster.chief() and cluster.worker_count == 1 and cluster.worker_index == 0
    strategy = cluster.execution_strategy()
    assert strategy.num_replicas_in_sync == 1
    with strategy.scope() as s:
        assert s is strategy


def test_worker_count_must_match_the_launch(monkeypatch):
    monkeypatch.setenv("RANK", "0"); monkeypatch.setenv("WORLD_SIZE", "1")
    with pytest.raises(b200xf.B200XFError):
        b200xf.B200Cluster(2)
    assert b200xf.B200Cluster(1).chief()


def _cluster_worker(out_dir):
    import b200xf as bx
    cluster = bx.B200Cluster(2, backend="gloo")
    strategy = cluster.execution_strategy()
    total = strategy.reduce(bx.ReduceOp.SUM, torch.tensor([float(cluster.worker_index + 1)]), axis=None)
    Path(out_dir, f"rank{cluster.worker_index}.txt").write_text(f"{int(cluster.chief())} {strategy.num_replicas_in_sync} {float(total)}")
    cluster.shutdown()


def test_launch_two_workers_rendezvous_and_chief(tmp_path):
    b200xf.launch(2, _cluster_worker, (str(tmp_path),))
    r0 = (tmp_path / "rank0.txt").read_text().split(); r1 = (tmp_path / "rank1.txt").read_text().split()
    assert r0 == ["1", "2", "3.0"] and r1 == ["0", "2", "3.0"]  # one chief, two replicas, SUM over both


def _failing_worker():
    raise RuntimeError("worker failure must reach the launcher")


def test_launch_propagates_worker_failure():
    with pytest.raises(Exception):
        b200xf.launch(2, _failing_worker)
    assert b200xf.launch(1, lambda: 7) == 7
