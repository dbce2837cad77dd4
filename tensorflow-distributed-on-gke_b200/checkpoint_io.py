"""TensorFlow tensor-bundle checkpoints under the reference's `save_weights(prefix)` / `load_weights(prefix)` call shape
(checkpoint.py:86 writes `<dir>/model_weights`, __main__.py:87 reads `saved_weights/2/model_weights`).

File set, as in /root/reference/saved_weights/<n>/: `<prefix>.index` (LevelDB-style SSTable: prefix-compressed keys,
uncompressed blocks with masked-CRC32C trailers, 48-byte footer, magic 0xdb4775248b80fb57; key "" -> BundleHeaderProto,
every other key -> BundleEntryProto{dtype, shape, shard_id, offset, size, crc32c}), `<prefix>.data-00000-of-00001` (the
tensors, little-endian, densely packed in ascending offset) and `checkpoint` (two text lines).  Keys are the Keras
object-graph names `<variable>/.ATTRIBUTES/VARIABLE_VALUE`; the data shard is laid out in the reference's physical order
(final_layer, embeddings, all LayerNorms, encoder layers, decoder layers), so for the tutorial model every offset and
size equals the reference's own index (tests/test_checkpoint_io.py checks this against tests/golden/param_table.json).

The `_CHECKPOINTABLE_OBJECT_GRAPH` entry (a DT_STRING scalar holding a serialized TrackableObjectGraph, which Keras'
object-based `load_weights` walks to find each variable) is written too, rebuilt from the reference's class structure
(transformer_model.py:112-122,169-191,207-225,251-269,282-300,318-333): nodes in breadth-first order from the model, children
in attribute-assignment order, Sequential children as layer_with_weights-i / layer-i pairs, one list node per
encoder_layers / decoder_layers and per LayerNormalization `axis`, Dropout layers as empty nodes, one node per variable with
the VARIABLE_VALUE attribute and its checkpoint key.  For the reference's tutorial model (4 layers) the result is 22 436 bytes
with masked CRC32C 0x71f2074f — exactly the size and checksum the reference's own saved_weights/{1,2}/model_weights.index
record for this entry (tests/test_checkpoint_io.py), although the reference's data shards themselves are missing.
"""
from __future__ import annotations

import struct
from pathlib import Path

import numpy as np

SUFFIX = "/.ATTRIBUTES/VARIABLE_VALUE"
MAGIC = 0xdb4775248b80fb57
DT_FLOAT = 1
DT_STRING = 7
OBJECT_GRAPH_KEY = "_CHECKPOINTABLE_OBJECT_GRAPH"
BLOCK_TARGET = 4096       # leveldb table block size
RESTART_INTERVAL = 16

# ---------------------------------------------------------------------------------------------- crc32c (Castagnoli)
_TABLE = None


def _table():
    global _TABLE
    if _TABLE is None:
        t = np.zeros((8, 256), dtype=np.uint32)
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ (0x82F63B78 if c & 1 else 0)
            t[0, i] = c
        for k in range(1, 8):
            t[k] = (t[k - 1] >> 8) ^ t[0][t[k - 1] & 0xFF]
        _TABLE = t
    return _TABLE


def crc32c(data: bytes, crc: int = 0) -> int:
    """CRC-32C of `data`.  Uses the library's native routine when libb200xf.so is loadable, else a table walk."""
    try:
        from . import _lib
        return int(_lib.load().b200xf_crc32c(bytes(data), len(data), crc))
    except Exception:   # library not built / not loadable: the table walk below gives the same value
        pass
    t = _table()[0]
    c = crc ^ 0xFFFFFFFF
    for b in bytes(data):
        c = int(t[(c ^ b) & 0xFF]) ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc32c(data: bytes) -> int:
    """leveldb / TF mask: rotate right by 15 and add a constant."""
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xa282ead8) & 0xFFFFFFFF


# ---------------------------------------------------------------------------------------------- varints / protobuf
def _varint(v: int) -> bytes:
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(b, p):
    r = s = 0
    while True:
        c = b[p]; p += 1
        r |= (c & 0x7F) << s; s += 7
        if c < 0x80:
            return r, p


def _field(num, wire, payload: bytes) -> bytes:
    return _varint((num << 3) | wire) + payload


def _bundle_header() -> bytes:
    # BundleHeaderProto{num_shards=1, endianness=LITTLE(0, default, omitted), version{producer=1}}
    return _field(1, 0, _varint(1)) + _field(3, 2, _varint(2) + _field(1, 0, _varint(1)))


def _bundle_entry(shape, offset, size, crc) -> bytes:
    dims = b"".join(_field(2, 2, _varint(len(d)) + d) for d in (_field(1, 0, _varint(int(s))) for s in shape))
    out = _field(1, 0, _varint(DT_FLOAT)) + _field(2, 2, _varint(len(dims)) + dims)
    if offset:
        out += _field(4, 0, _varint(offset))
    out += _field(5, 0, _varint(size)) + _field(6, 5, struct.pack("<I", crc))
    return out


def _bundle_entry_string(offset, size, crc) -> bytes:
    # BundleEntryProto of a scalar DT_STRING tensor: dtype 7, an empty TensorShapeProto, shard 0 (omitted), offset, size, crc32c
    return _field(1, 0, _varint(DT_STRING)) + _field(2, 2, _varint(0)) + _field(4, 0, _varint(offset)) + _field(5, 0, _varint(size)) + \
        _field(6, 5, struct.pack("<I", crc))


def _lenf(num, payload: bytes) -> bytes:
    return _field(num, 2, _varint(len(payload)) + payload)


def object_graph_proto(layers_count: int) -> bytes:
    """Serialized TrackableObjectGraph of the reference's `Transformer` (keras.Model) with `layers_count` encoder and decoder
    layers, as `model.save_weights` (checkpoint.py:86) stores it under `_CHECKPOINTABLE_OBJECT_GRAPH`.

    Node = message {repeated children {node_id = 1, local_name = 2}; repeated attributes {name = 1, checkpoint_key = 3}}.
    Structure: see the module docstring; byte-exact against the reference's index for layers_count = 4."""
    class N:
        __slots__ = ("children", "is_var")

        def __init__(self, is_var=False):
            self.children = []; self.is_var = is_var

        def add(self, name, child):
            self.children.append((name, child)); return child

    def variables(node, *names):
        for n in names:
            node.add(n, N(True))
        return node

    def dense():
        return variables(N(), "kernel", "bias")

    def layernorm():
        n = N(); n.add("axis", N())          # LayerNormalization.build turns `axis` into a (tracked, empty) list
        return variables(n, "gamma", "beta")

    def mha():
        n = N()
        for name in ("query_generator_weights", "key_generator_weights", "value_generator_weights", "dense"):
            n.add(name, dense())
        return n

    def ffn():                               # keras.Sequential: layer_with_weights-i and layer-i name the same child
        n = N()
        for i in range(2):
            d = dense(); n.add(f"layer_with_weights-{i}", d); n.add(f"layer-{i}", d)
        return n

    def stack(list_name, make_layer):
        n = N(); variables(n.add("embedding", N()), "embeddings")
        lst = n.add(list_name, N())
        for i in range(layers_count):
            lst.add(str(i), make_layer())
        n.add("dropout", N())
        return n

    def encoder_layer():
        n = N(); n.add("mha", mha()); n.add("ffn", ffn())
        for k in (1, 2): n.add(f"layernorm{k}", layernorm())
        for k in (1, 2): n.add(f"dropout{k}", N())
        return n

    def decoder_layer():
        n = N(); n.add("mha1", mha()); n.add("mha2", mha()); n.add("ffn", ffn())
        for k in (1, 2, 3): n.add(f"layernorm{k}", layernorm())
        for k in (1, 2, 3): n.add(f"dropout{k}", N())
        return n

    root = N()
    root.add("encoder", stack("encoder_layers", encoder_layer)); root.add("decoder", stack("decoder_layers", decoder_layer))
    root.add("final_layer", dense())
    # breadth-first numbering; a node's checkpoint path is the path it was first reached by
    order = [root]; ids = {id(root): 0}; path = {id(root): ""}
    head = 0
    while head < len(order):
        n = order[head]; head += 1
        for name, c in n.children:
            if id(c) not in ids:
                ids[id(c)] = len(order); order.append(c)
                path[id(c)] = (path[id(n)] + "/" if path[id(n)] else "") + name
    out = bytearray()
    for n in order:
        body = bytearray()
        for name, c in n.children:
            body += _lenf(1, _field(1, 0, _varint(ids[id(c)])) + _lenf(2, name.encode()))
        if n.is_var:
            body += _lenf(2, _lenf(1, b"VARIABLE_VALUE") + _lenf(3, (path[id(n)] + SUFFIX).encode()))
        out += _lenf(1, bytes(body))
    return bytes(out)


def _mask(c: int) -> int:
    return (((c >> 15) | (c << 17)) + 0xa282ead8) & 0xFFFFFFFF


def string_tensor_bytes(value: bytes):
    """On-disk form of a scalar DT_STRING tensor in a bundle's data shard, and the checksum its index entry carries:
    [varint length][masked CRC32C of the length as a little-endian uint32][the bytes]; the entry's crc32c is the masked
    CRC32C over the uint32 length, that 4-byte length checksum and the bytes."""
    c = crc32c(struct.pack("<I", len(value)))
    length_checksum = struct.pack("<I", _mask(c))
    c = crc32c(value, crc32c(length_checksum, c))
    return _varint(len(value)) + length_checksum + value, _mask(c)


def _parse_proto(b):
    p = 0; f = {}
    while p < len(b):
        tag, p = _read_varint(b, p); num, wire = tag >> 3, tag & 7
        if wire == 0: v, p = _read_varint(b, p)
        elif wire == 2:
            n, p = _read_varint(b, p); v = b[p:p + n]; p += n
        elif wire == 5: v = struct.unpack("<I", b[p:p + 4])[0]; p += 4
        elif wire == 1: v = struct.unpack("<Q", b[p:p + 8])[0]; p += 8
        else: raise ValueError(f"unsupported wire type {wire}")
        f.setdefault(num, []).append(v)
    return f


# ---------------------------------------------------------------------------------------------- SSTable
class _BlockBuilder:
    def __init__(self):
        self.buf = bytearray(); self.restarts = [0]; self.count = 0; self.last = b""

    def add(self, key: bytes, value: bytes):
        shared = 0
        if self.count % RESTART_INTERVAL == 0 and self.count:
            self.restarts.append(len(self.buf))
        elif self.count:
            n = min(len(key), len(self.last))
            while shared < n and key[shared] == self.last[shared]:
                shared += 1
        self.buf += _varint(shared) + _varint(len(key) - shared) + _varint(len(value)) + key[shared:] + value
        self.last = key; self.count += 1

    def size(self):
        return len(self.buf) + 4 * len(self.restarts) + 4

    def finish(self) -> bytes:
        return bytes(self.buf) + b"".join(struct.pack("<I", r) for r in self.restarts) + struct.pack("<I", len(self.restarts))


def _write_table(path: Path, items):
    """items: sorted [(key bytes, value bytes)].  Uncompressed blocks, each followed by type byte 0 + masked crc32c."""
    out = bytearray(); index = _BlockBuilder()

    def emit(block: bytes):
        off = len(out)
        out.extend(block); out.extend(b"\x00"); out.extend(struct.pack("<I", masked_crc32c(block + b"\x00")))
        return off, len(block)

    cur = _BlockBuilder(); last_key = b""
    for k, v in items:
        cur.add(k, v); last_key = k
        if cur.size() >= BLOCK_TARGET:
            off, n = emit(cur.finish()); index.add(last_key, _varint(off) + _varint(n)); cur = _BlockBuilder()
    if cur.count:
        off, n = emit(cur.finish()); index.add(last_key, _varint(off) + _varint(n))
    moff, mn = emit(_BlockBuilder().finish())          # empty metaindex block
    ioff, inn = emit(index.finish())
    footer = _varint(moff) + _varint(mn) + _varint(ioff) + _varint(inn)
    footer += b"\x00" * (40 - len(footer)) + struct.pack("<Q", MAGIC)
    out.extend(footer)
    path.write_bytes(bytes(out))


def _read_block(d, off, size, verify=True):
    blk = d[off:off + size]
    if verify:
        want = struct.unpack("<I", d[off + size + 1:off + size + 5])[0]
        if masked_crc32c(d[off:off + size + 1]) != want:
            raise ValueError("index block checksum mismatch")
    nrest = struct.unpack("<I", blk[-4:])[0]
    end = len(blk) - 4 - 4 * nrest
    p = 0; key = b""; out = []
    while p < end:
        shared, p = _read_varint(blk, p); non, p = _read_varint(blk, p); vlen, p = _read_varint(blk, p)
        key = key[:shared] + blk[p:p + non]; p += non
        out.append((key, blk[p:p + vlen])); p += vlen
    return out


def read_index(path):
    """Decode `<prefix>.index` -> list of dicts (key, dtype, shape, offset, size, crc32c), header entry excluded."""
    d = Path(path).read_bytes()
    foot = d[-48:]
    if struct.unpack("<Q", foot[-8:])[0] != MAGIC:
        raise ValueError(f"{path}: not a tensor-bundle index (bad magic)")
    p = 0
    _, p = _read_varint(foot, p); _, p = _read_varint(foot, p)
    ioff, p = _read_varint(foot, p); isz, p = _read_varint(foot, p)
    out = []
    for _, handle in _read_block(d, ioff, isz):
        q = 0; boff, q = _read_varint(handle, q); bsz, q = _read_varint(handle, q)
        for k, v in _read_block(d, boff, bsz):
            if k == b"":
                continue
            f = _parse_proto(v)
            shape = [(_parse_proto(dim).get(1, [0])[0]) for dim in _parse_proto(f[2][0]).get(2, [])] if 2 in f else []
            out.append(dict(key=k.decode(), dtype=f.get(1, [0])[0], shape=shape, offset=f.get(4, [0])[0], size=f.get(5, [0])[0],
                            crc32c=f.get(6, [0])[0]))
    return out


# ---------------------------------------------------------------------------------------------- public API
def reference_order(names):
    """The reference's physical data-shard order: final_layer, embeddings, every LayerNorm, then the
    encoder layers' and the decoder layers' Dense variables — which is the order `Transformer.trainable_variables` uses."""
    return list(names)


def save_weights(model, prefix):
    prefix = Path(str(prefix)); prefix.parent.mkdir(parents=True, exist_ok=True)
    weights = model.get_weights_dict()                      # already in the reference's physical order
    entries = {}; offset = 0
    with open(str(prefix) + ".data-00000-of-00001", "wb") as f:
        for name, a in weights.items():
            raw = np.ascontiguousarray(a, dtype="<f4").tobytes()
            f.write(raw)
            entries[(name + SUFFIX).encode()] = _bundle_entry(a.shape, offset, len(raw), masked_crc32c(raw))
            offset += len(raw)
    # the Keras object graph, so that the reference's `transformer.load_weights(prefix)` (__main__.py:87) can restore from this file
    layers_count = 1 + max(int(n.split("/")[2]) for n in weights if n.startswith("encoder/encoder_layers/"))
    with open(str(prefix) + ".data-00000-of-00001", "ab") as f:
        blob, crc = string_tensor_bytes(object_graph_proto(layers_count))
        f.write(blob)
    entries[OBJECT_GRAPH_KEY.encode()] = _bundle_entry_string(offset, len(blob), crc)
    items = [(b"", _bundle_header())] + sorted(entries.items())
    _write_table(Path(str(prefix) + ".index"), items)
    (prefix.parent / "checkpoint").write_text(f'model_checkpoint_path: "{prefix.name}"\nall_model_checkpoint_paths: "{prefix.name}"\n')


def load_arrays(prefix, verify=True):
    """{variable name: float32 ndarray} from a tensor bundle (the reference's own, or one written by save_weights)."""
    prefix = str(prefix)
    data_path = Path(prefix + ".data-00000-of-00001")
    if not data_path.exists():
        raise FileNotFoundError(f"{data_path}: data shard missing (the reference repository ships only the .index files)")
    out = {}
    with open(data_path, "rb") as f:
        for e in read_index(prefix + ".index"):
            if not e["key"].endswith(SUFFIX) or e["dtype"] != DT_FLOAT:
                continue                                      # e.g. _CHECKPOINTABLE_OBJECT_GRAPH (DT_STRING)
            f.seek(e["offset"]); raw = f.read(e["size"])
            if len(raw) != e["size"]:
                raise ValueError(f"{e['key']}: data shard truncated")
            if verify and masked_crc32c(raw) != e["crc32c"]:
                raise ValueError(f"{e['key']}: tensor checksum mismatch")
            out[e["key"][:-len(SUFFIX)]] = np.frombuffer(raw, dtype="<f4").reshape(e["shape"]).copy()
    return out


def load_weights(model, prefix):
    arrays = load_arrays(prefix)
    have = {v.name: v.shape for v in model.trainable_variables}
    missing = sorted(set(have) - set(arrays))
    if missing:
        raise KeyError(f"checkpoint lacks {len(missing)} variables, e.g. {missing[:3]}")
    for n, shp in have.items():
        if tuple(arrays[n].shape) != tuple(shp):
            raise ValueError(f"{n}: checkpoint shape {arrays[n].shape} != model shape {shp}")
    model.set_weights_dict(arrays)
