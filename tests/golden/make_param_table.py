"""Decode the reference's TF tensor-bundle index into a JSON fixture.

Run in the build container only (reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_param_table.py
Source artefact: /root/reference/saved_weights/2/model_weights.index (written by
checkpoint.py:86 `model.save_weights`, read back by __main__.py:87 `load_weights`).
Format: LevelDB-style SSTable (prefix-compressed keys, 48-byte footer) whose values are
BundleEntryProto messages {1:dtype, 2:shape{2:dim{1:size}}, 3:shard_id, 4:offset, 5:size, 6:crc32c}.
"""
import json, struct, sys
from pathlib import Path

SRC = Path("/root/reference/saved_weights/2/model_weights.index")
SRC1 = Path("/root/reference/saved_weights/1/model_weights.index")
OUT = Path(__file__).with_name("param_table.json")
MAGIC = 0xdb4775248b80fb57


def varint(b, p):
    r = s = 0
    while True:
        c = b[p]; p += 1
        r |= (c & 0x7F) << s; s += 7
        if c < 0x80:
            return r, p


def read_block(d, off, size):
    blk = d[off:off + size]
    nrest = struct.unpack("<I", blk[-4:])[0]
    end = len(blk) - 4 - 4 * nrest
    p = 0; key = b""; out = []
    while p < end:
        shared, p = varint(blk, p); non, p = varint(blk, p); vlen, p = varint(blk, p)
        key = key[:shared] + blk[p:p + non]; p += non
        out.append((key, blk[p:p + vlen])); p += vlen
    return out


def parse_proto(b):
    p = 0; f = {}
    while p < len(b):
        tag, p = varint(b, p); fn, wt = tag >> 3, tag & 7
        if wt == 0: v, p = varint(b, p)
        elif wt == 2:
            l, p = varint(b, p); v = b[p:p + l]; p += l
        elif wt == 5: v = struct.unpack("<I", b[p:p + 4])[0]; p += 4
        elif wt == 1: v = struct.unpack("<Q", b[p:p + 8])[0]; p += 8
        else: raise ValueError(wt)
        f.setdefault(fn, []).append(v)
    return f


def decode(path):
    d = path.read_bytes()
    foot = d[-48:]
    assert struct.unpack("<Q", foot[-8:])[0] == MAGIC
    p = 0
    _, p = varint(foot, p); _, p = varint(foot, p)
    ioff, p = varint(foot, p); isz, p = varint(foot, p)
    entries = []
    for _, handle in read_block(d, ioff, isz):
        q = 0; boff, q = varint(handle, q); bsz, q = varint(handle, q)
        entries += read_block(d, boff, bsz)
    out = []
    for k, v in entries:
        if k == b"": continue
        f = parse_proto(v)
        shape = []
        if 2 in f:
            for dim in parse_proto(f[2][0]).get(2, []):
                shape.append(parse_proto(dim).get(1, [0])[0])
        out.append(dict(key=k.decode(), dtype=f.get(1, [0])[0], shape=shape,
                        offset=f.get(4, [0])[0], size=f.get(5, [0])[0], crc32c=f.get(6, [0])[0]))
    return out


if __name__ == "__main__":
    e2 = decode(SRC); e1 = decode(SRC1)
    assert [(e["key"], e["shape"], e["offset"]) for e in e1] == [(e["key"], e["shape"], e["offset"]) for e in e2]
    suffix = "/.ATTRIBUTES/VARIABLE_VALUE"
    tensors = [dict(name=e["key"][:-len(suffix)], shape=e["shape"], offset=e["offset"], size=e["size"])
               for e in e2 if e["dtype"] == 1]
    tensors.sort(key=lambda t: t["offset"])
    total = sum(t["size"] for t in tensors)
    # the Keras object graph (DT_STRING scalar): both checkpoints record the same offset, size and masked CRC32C for it
    og2 = [e for e in e2 if e["key"] == "_CHECKPOINTABLE_OBJECT_GRAPH"][0]; og1 = [e for e in e1 if e["key"] == "_CHECKPOINTABLE_OBJECT_GRAPH"][0]
    assert og1 == og2
    OUT.write_text(json.dumps(dict(source=str(SRC), n_tensors=len(tensors), total_bytes=total,
                                   total_params=total // 4, object_graph=og2, tensors=tensors), indent=1))
    print(len(tensors), total, total // 4)
    for t in tensors[:30]: print(t)
