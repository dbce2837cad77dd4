"""TF tensor-bundle checkpoint files: the writer reproduces the reference's own index layout, the
reader decodes the reference's real `.index` files, and a write -> read round trip is exact.  CPU only."""
import json
from collections import OrderedDict
from pathlib import Path

import numpy as np
import pytest

import b200xf
from b200xf import checkpoint_io as ck
from oracle import transformer_oracle as O

GOLD = json.loads((Path(__file__).parent / "golden" / "param_table.json").read_text())
REF = Path("/root/reference/saved_weights")


class _HostModel:
    """Stands in for Transformer on a box without a GPU: same variable names / shapes / order."""

    def __init__(self, arrays):
        self.arrays = OrderedDict(arrays)

    def get_weights_dict(self):
        return self.arrays

    @property
    def trainable_variables(self):
        class V:  # noqa
            def __init__(s, n, a): s.name, s.shape = n, a.shape
        return [V(n, a) for n, a in self.arrays.items()]

    def set_weights_dict(self, arrays):
        self.arrays = OrderedDict((n, np.asarray(arrays[n])) for n in self.arrays)


def test_crc32c_known_answers():
    assert ck.crc32c(b"123456789") == 0xE3069283          # the standard CRC-32C check value
    assert ck.crc32c(b"") == 0
    assert ck.crc32c(b"6789", ck.crc32c(b"12345")) == 0xE3069283
    # pure-Python table walk and the native routine agree
    t = ck._table()[0]; c = 0xFFFFFFFF
    for b in b"tensor bundle":
        c = int(t[(c ^ b) & 0xFF]) ^ (c >> 8)
    assert (c ^ 0xFFFFFFFF) == ck.crc32c(b"tensor bundle")


def test_written_index_matches_the_reference_layout(tmp_path):
    arrays = O.init_parameters(4, 128, 512, 7765, 7010, seed=3)     # the tutorial model: 172 variables, 18 587 528 bytes
    ck.save_weights(_HostModel(arrays), tmp_path / "model_weights")
    got = {e["key"]: e for e in ck.read_index(str(tmp_path / "model_weights.index"))}
    og = got.pop(ck.OBJECT_GRAPH_KEY)
    # The Keras object graph that load_weights (__main__.py:87) restores by: the reference's own index records this entry's
    # offset, size and masked CRC32C (its data shard is missing), and the rebuilt graph reproduces all three exactly.
    assert {k: og[k] for k in ("dtype", "shape", "offset", "size", "crc32c")} == {k: GOLD["object_graph"][k] for k in ("dtype", "shape", "offset", "size", "crc32c")}
    assert (og["size"], og["crc32c"]) == (22436, 1911686991)
    assert len(got) == 172
    for t in GOLD["tensors"]:                                        # offsets, sizes and shapes of the reference's own index
        e = got[t["name"] + ck.SUFFIX]
        assert (e["offset"], e["size"], e["shape"], e["dtype"]) == (t["offset"], t["size"], t["shape"], 1), t["name"]
    assert GOLD["total_bytes"] == 18587528
    assert (tmp_path / "model_weights.data-00000-of-00001").stat().st_size == GOLD["total_bytes"] + 22436   # tensors + object graph
    assert (tmp_path / "checkpoint").read_text() == 'model_checkpoint_path: "model_weights"\nall_model_checkpoint_paths: "model_weights"\n'
    back = ck.load_arrays(tmp_path / "model_weights")
    assert list(back) == sorted(arrays)                              # SSTable keys are sorted
    assert all(np.array_equal(back[n], arrays[n]) for n in arrays)
    m = _HostModel({n: np.zeros_like(a) for n, a in arrays.items()})
    ck.load_weights(m, tmp_path / "model_weights")
    assert all(np.array_equal(m.arrays[n], arrays[n]) for n in arrays)


def test_corruption_is_detected(tmp_path):
    arrays = O.init_parameters(1, 32, 64, 50, 40, seed=1)
    ck.save_weights(_HostModel(arrays), tmp_path / "w")
    data = tmp_path / "w.data-00000-of-00001"
    raw = bytearray(data.read_bytes()); raw[100] ^= 0xFF; data.write_bytes(bytes(raw))
    with pytest.raises(ValueError, match="checksum"):
        ck.load_arrays(tmp_path / "w")
    with pytest.raises(ValueError):
        m = _HostModel({n: np.zeros((3,) + a.shape) for n, a in arrays.items()})
        data.write_bytes(b"".join(np.ascontiguousarray(a, "<f4").tobytes() for a in arrays.values()))
        ck.load_weights(m, tmp_path / "w")


@pytest.mark.skipif(not REF.exists(), reason="/root/reference is only present in the build container")
def test_reader_decodes_the_reference_index_files():
    for n in ("1", "2"):
        entries = ck.read_index(str(REF / n / "model_weights.index"))    # verifies every block checksum on the way
        floats = [e for e in entries if e["dtype"] == 1]
        assert len(floats) == 172 and len(entries) == 173                  # + _CHECKPOINTABLE_OBJECT_GRAPH (DT_STRING)
        by_key = {e["key"]: e for e in floats}
        for t in GOLD["tensors"]:
            e = by_key[t["name"] + ck.SUFFIX]
            assert (e["offset"], e["size"], e["shape"]) == (t["offset"], t["size"], t["shape"])
        with pytest.raises(FileNotFoundError):                             # the data shards are not in the repository
            ck.load_arrays(REF / n / "model_weights")


def test_object_graph_structure():
    """The TrackableObjectGraph written under _CHECKPOINTABLE_OBJECT_GRAPH: every variable's checkpoint key appears exactly once
    as an attribute, node ids are a breadth-first numbering, and the graph scales with the layer count."""
    for L in (1, 4, 6):
        proto = ck.object_graph_proto(L)
        nodes = [ck._parse_proto(n) for n in ck._parse_proto(proto)[1]]
        keys = []
        for n in nodes:
            for att in n.get(2, []):
                f = ck._parse_proto(att)
                assert f[1] == [b"VARIABLE_VALUE"] and 2 not in f       # name, no full_name (as the reference's TF wrote it)
                keys.append(f[3][0].decode())
        names = O.parameter_names(L)
        assert sorted(keys) == sorted(n + ck.SUFFIX for n in names) and len(keys) == len(set(keys))
        ids = [ck._parse_proto(c)[1][0] for n in nodes for c in n.get(1, [])]
        assert set(ids) == set(range(1, len(nodes)))                      # every node but the root is somebody's child
        # root: encoder, decoder, final_layer
        assert [ck._parse_proto(c)[2][0] for c in nodes[0][1]] == [b"encoder", b"decoder", b"final_layer"]
    assert len(ck.object_graph_proto(4)) == 22429
    blob, crc = ck.string_tensor_bytes(ck.object_graph_proto(4))
    assert len(blob) == 22436 and crc == 0x71f2074f                        # saved_weights/{1,2}/model_weights.index of the reference
