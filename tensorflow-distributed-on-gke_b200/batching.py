"""On-device batch assembly.

The reference tokenises a *global* batch of sentences and pads it to its longest sentence with 0 = PAD
(`english_portugese_dataset.py:36-39`, `.to_tensor()`), then `experimental_distribute_dataset` gives every replica a
contiguous slice of the rows (`__main__.py:48-53`, AutoShardPolicy.DATA `english_portugese_dataset.py:27-31`).  Every new
(batch, length) pair makes `tf.function` retrace (`__main__.py:95-99,127`).

Here the ragged sentences travel to the GPU back to back (fewer bytes than the padded matrix), the padding and the
per-replica split happen in one kernel (`b200xf_assemble_padded`), and the padded length can be rounded up to a multiple
(`length_multiple`) so that the number of distinct (B, S, T) shapes -- and with it the number of captured CUDA graphs --
stays bounded.  Rounding up only appends PAD columns, which neither the loss nor the attention masks see.

The host part (lengths, bucketing, slice bounds) is plain numpy and is what the CPU tests cover; the kernel is covered by
the GPU tests against `pad_to_longest` below.
"""
import numpy as np

from . import _lib
from ._lib import check, ptr

PAD = 0


def ragged_to_flat(sentences):
    """list of 1-D integer sequences -> (flat int64 [total], offsets int64 [n + 1])."""
    lengths = np.fromiter((len(s) for s in sentences), dtype=np.int64, count=len(sentences))
    offsets = np.zeros(len(sentences) + 1, dtype=np.int64)
    np.cumsum(lengths, out=offsets[1:])
    flat = np.empty(int(offsets[-1]), dtype=np.int64)
    for s, b in zip(sentences, offsets[:-1]):
        flat[b:b + len(s)] = np.asarray(s, dtype=np.int64)
    return flat, offsets


def padded_length(offsets, length_multiple=1, maximum_length=None):
    """Length every row of the GLOBAL batch is padded to: the longest sentence, rounded up to `length_multiple`."""
    if length_multiple < 1:
        raise ValueError("length_multiple must be >= 1")
    lengths = np.diff(np.asarray(offsets, dtype=np.int64))
    longest = int(lengths.max()) if lengths.size else 0
    if longest <= 0:
        raise ValueError("batch without tokens")
    L = -(-longest // length_multiple) * length_multiple
    if maximum_length is not None:
        if longest > maximum_length:
            raise ValueError(f"sentence of {longest} tokens exceeds the maximum length {maximum_length}")
        L = min(L, int(maximum_length))
    return L


def replica_rows(global_rows, rank, world):
    """Contiguous slice (first row, row count) of a global batch that replica `rank` keeps — the rule of
    B200Strategy.experimental_distribute_dataset.  No row is dropped: when the global batch does not divide evenly (the
    last batch of an epoch, __main__.py:95-99) the first `rows % world` replicas take one row more, as the per-replica
    re-batching of tf.distribute does (sizes differ by at most one)."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside a world of {world}")
    n, rem = divmod(global_rows, world)
    return rank * n + min(rank, rem), n + (1 if rank < rem else 0)


def pad_to_longest(sentences, length=None):
    """numpy statement of `.to_tensor()`: [n, length] int64, PAD = 0 (test oracle for the kernel)."""
    L = max(len(s) for s in sentences) if length is None else length
    out = np.zeros((len(sentences), L), dtype=np.int64)
    for r, s in enumerate(sentences):
        out[r, :min(len(s), L)] = np.asarray(s[:L], dtype=np.int64)
    return out


class BatchAssembler:
    """assembler = BatchAssembler(rank, world, length_multiple=16, maximum_input_length=..., maximum_target_length=...)
    inp, tgt = assembler.assemble(portuguese_token_lists, english_token_lists)   # device int64 [B / world, S], [B / world, T + 1]

    Buffers are cached per shape, so a steady stream of batches allocates nothing; one H2D copy per side and step."""

    def __init__(self, rank=0, world=1, length_multiple=1, maximum_input_length=None, maximum_target_length=None, device=None):
        import torch
        if not torch.cuda.is_available():
            raise _lib.B200XFError("BatchAssembler needs a CUDA device (there is no CPU fallback)")
        self.rank, self.world, self.length_multiple = int(rank), int(world), int(length_multiple)
        self.maximum_input_length, self.maximum_target_length = maximum_input_length, maximum_target_length
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self._lib = _lib.load()
        self._out = {}
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.shapes_seen = set()

    def _side(self, sentences, maximum_length, tag):
        import torch
        flat, offsets = ragged_to_flat(sentences)
        L = padded_length(offsets, self.length_multiple, maximum_length)
        first, rows = replica_rows(len(sentences), self.rank, self.world)
        if rows == 0:
            raise ValueError(f"global batch of {len(sentences)} sentences is smaller than the world size {self.world}")
        # only this replica's sentences cross the bus
        lo, hi = int(offsets[first]), int(offsets[first + rows])
        flat_d = torch.from_numpy(flat[lo:hi]).to(self.device, non_blocking=True)
        off_d = torch.from_numpy(offsets[first:first + rows + 1] - lo).to(self.device, non_blocking=True)
        key = (tag, rows, L)
        out = self._out.get(key)
        if out is None:
            out = self._out[key] = torch.empty((rows, L), dtype=torch.int64, device=self.device)
        check(self._lib.b200xf_assemble_padded(ptr(flat_d), ptr(off_d), 0, rows, L, ptr(out), ptr(self._flag), _lib.stream_ptr()))
        return out

    def assemble(self, input_sentences, target_sentences):
        if len(input_sentences) != len(target_sentences):
            raise ValueError("inputs and targets differ in number of sentences")
        inp = self._side(input_sentences, self.maximum_input_length, "inp")
        tgt = self._side(target_sentences, self.maximum_target_length, "tgt")
        self.shapes_seen.add((inp.shape[0], inp.shape[1], tgt.shape[1]))
        return inp, tgt

    def check_lengths(self):
        """Synchronises; raises if any sentence since the last check was longer than its row (cannot happen through
        `assemble`, which sizes the rows from the lengths; kept for callers that pass their own length)."""
        n = int(self._flag.item())
        self._flag.zero_()
        if n:
            raise _lib.B200XFError(f"a sentence of {n} tokens was truncated")
