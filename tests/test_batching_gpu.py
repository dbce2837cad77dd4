"""GPU tests of on-device batch assembly: `b200xf_assemble_padded` / `BatchAssembler` against the
numpy statement of the reference's `.to_tensor()` padding and per-replica split (english_portugese_dataset.py:36-47,
__main__.py:48-53), and a train step fed from it against the same step fed from host-padded arrays."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import b200xf  # noqa: E402
from b200xf._lib import check, load, ptr, stream_ptr  # noqa: E402

pytestmark = pytest.mark.gpu


def _ragged(n, lo, hi, vocab, seed):
    rng = np.random.default_rng(seed)
    return [[2] + list(rng.integers(4, vocab, size=int(rng.integers(lo, hi + 1)))) + [3] for _ in range(n)]


@pytest.mark.parametrize("rank,world", [(0, 1), (1, 3), (3, 4)])
@pytest.mark.parametrize("multiple", [1, 16])
def test_assembler_matches_pad_to_longest(rank, world, multiple):
    inp, tgt = _ragged(24, 1, 40, 200, 3), _ragged(24, 1, 33, 150, 4)
    asm = b200xf.BatchAssembler(rank, world, length_multiple=multiple)
    d_inp, d_tgt = asm.assemble(inp, tgt)
    asm.check_lengths()
    first, rows = b200xf.replica_rows(24, rank, world)
    for dev, sents in ((d_inp, inp), (d_tgt, tgt)):
        L = b200xf.padded_length(b200xf.ragged_to_flat(sents)[1], multiple)   # the GLOBAL batch's longest sentence
        want = b200xf.pad_to_longest(sents, L)[first:first + rows]
        assert dev.shape == want.shape and dev.dtype == torch.int64
        assert np.array_equal(dev.cpu().numpy(), want)
    # the same shapes again reuse the same buffers
    again = asm.assemble(inp, tgt)
    assert again[0].data_ptr() == d_inp.data_ptr() and len(asm.shapes_seen) == 1


def test_too_short_rows_are_reported():
    sents = [[5, 6, 7, 8, 9], [4]]
    flat, off = b200xf.ragged_to_flat(sents)
    d_flat, d_off = torch.from_numpy(flat).cuda(), torch.from_numpy(off).cuda()
    out = torch.full((2, 3), -1, dtype=torch.int64, device="cuda"); flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    check(load().b200xf_assemble_padded(ptr(d_flat), ptr(d_off), 0, 2, 3, ptr(out), ptr(flag), stream_ptr()))
    assert out.cpu().tolist() == [[5, 6, 7], [4, 0, 0]] and int(flag.item()) == 5
    assert load().b200xf_assemble_padded(ptr(d_flat), ptr(d_off), 0, 0, 3, ptr(out), ptr(flag), stream_ptr()) != 0  # empty batch


def test_train_step_from_assembled_batch_equals_host_padded_batch():
    hp = dict(layers_count=1, embedding_size=64, heads_count=2, feed_forward_hidden_size=128, input_vocabulary_size=200,
              target_vocabulary_size=150, maximum_input_length=64, maximum_target_length=64, droput_rate=0.0)
    inp, tgt = _ragged(8, 3, 20, 200, 5), _ragged(8, 3, 18, 150, 6)
    a = b200xf.Transformer(**hp, seed=3); b = b200xf.Transformer(**hp, seed=3)
    a.iterations = 50; b.iterations = 50
    asm = b200xf.BatchAssembler(0, 1, length_multiple=8, maximum_input_length=64, maximum_target_length=64)
    d_inp, d_tgt = asm.assemble(inp, tgt)
    ra = a.train_step_device(d_inp, d_tgt, 1).cpu().numpy()
    # host path: padded to the longest sentence only (no bucketing) -- extra PAD columns must not change anything but
    # the split-K / reduction order
    rb = b.train_step_device(b200xf.pad_to_longest(inp), b200xf.pad_to_longest(tgt), 1).cpu().numpy()
    np.testing.assert_allclose(ra, rb, rtol=2e-4)
    assert ra[2] == rb[2] == sum(len(s) - 1 for s in tgt)  # accuracy denominator = non-PAD target positions
