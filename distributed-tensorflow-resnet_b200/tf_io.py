"""Readers / writers for the reference's own TensorFlow artefacts, so that weights move between the
reference and this build in both directions without TensorFlow installed:

  * frozen GraphDef `.pb`   — `resnet_cifar_frozen_model.py` output, loaded by
                              `resnet_cifar_predict_from_pd.py:66-103` (`import_graph_def`); here:
                              `read_frozen_graph(path)` -> {TF variable name: float32 array} and
                              `write_frozen_graph(path, tensors)` (Const nodes, same naming).
  * TF-V2 checkpoint bundle — `model.ckpt-N.index` (a LevelDB-format sorted string table of
                              BundleEntryProto records) + `model.ckpt-N.data-SSSSS-of-NNNNN`
                              (raw little-endian tensor bytes), what `tf.train.Saver` /
                              MonitoredTrainingSession write (`resnet_cifar_train.py:315-319`) and
                              `saver.restore` reads (`resnet_cifar_eval.py:85`); here:
                              `read_checkpoint_index`, `read_checkpoint`, `write_checkpoint`.

The table / bundle formats are restated from their public specifications (LevelDB table_format.md;
tensorflow/core/protobuf/tensor_bundle.proto).  The reference's fixture
`test/resnet50-cifar-ckpt-20190218/model.ckpt-107738.index` (403 entries) pins the reader.
Protobuf messages are handled with the `tensorboard.compat.proto` stubs where they exist
(GraphDef, TensorProto, TensorShapeProto); BundleHeaderProto / BundleEntryProto are tiny and are
encoded / decoded by hand.
"""
import os
import struct
from collections import OrderedDict

import numpy as np

# tensorflow DataType enum values
DT_FLOAT, DT_DOUBLE, DT_INT32, DT_UINT8, DT_INT64, DT_BOOL = 1, 2, 3, 4, 9, 10
_NP = {DT_FLOAT: np.float32, DT_DOUBLE: np.float64, DT_INT32: np.int32, DT_UINT8: np.uint8, DT_INT64: np.int64,
       DT_BOOL: np.bool_}
_TF = {np.dtype(v): k for k, v in _NP.items()}

_TABLE_MAGIC = 0xdb4775248b80fb57
_MASK_DELTA = 0xa282ead8


# ------------------------------------------------------------------------------------------ crc32c
def crc32c(data, crc=0):
    """CRC-32C (Castagnoli).  Large buffers go through the library's host routine (rn_crc32c)."""
    from ._lib import lib
    import ctypes as C
    if isinstance(data, np.ndarray):
        buf = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
        return int(lib.rn_crc32c(buf.ctypes.data_as(C.c_void_p), buf.size, crc))
    b = bytes(data)
    return int(lib.rn_crc32c(C.c_char_p(b), len(b), crc))


def _mask(crc):
    return ((((crc >> 15) | (crc << 17)) & 0xffffffff) + _MASK_DELTA) & 0xffffffff


def _unmask(m):
    rot = (m - _MASK_DELTA) & 0xffffffff
    return ((rot >> 17) | (rot << 15)) & 0xffffffff


# ------------------------------------------------------------------------------------------ varints / protobuf
def _get_varint(b, p):
    r = 0
    s = 0
    while True:
        c = b[p]
        p += 1
        r |= (c & 0x7f) << s
        if c < 0x80:
            return r, p
        s += 7


def _put_varint(v):
    out = bytearray()
    v &= (1 << 64) - 1
    while v >= 0x80:
        out.append((v & 0x7f) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _pb_fields(b):
    """[(field number, wire type, value)] of one protobuf message (value: int or bytes)."""
    p, out = 0, []
    while p < len(b):
        key, p = _get_varint(b, p)
        f, w = key >> 3, key & 7
        if w == 0:
            v, p = _get_varint(b, p)
        elif w == 1:
            v = struct.unpack_from("<Q", b, p)[0]
            p += 8
        elif w == 2:
            n, p = _get_varint(b, p)
            v = bytes(b[p:p + n])
            p += n
        elif w == 5:
            v = struct.unpack_from("<I", b, p)[0]
            p += 4
        else:
            raise ValueError("unsupported protobuf wire type %d" % w)
        out.append((f, w, v))
    return out


def _parse_shape(b):
    dims = []
    for f, _, v in _pb_fields(b):
        if f == 2:                                  # repeated Dim dim = 2 { int64 size = 1; string name = 2 }
            size = 0
            for f2, _, v2 in _pb_fields(v):
                if f2 == 1:
                    size = v2 if v2 < (1 << 63) else v2 - (1 << 64)
            dims.append(size)
    return dims


def _encode_shape(shape):
    out = b""
    for d in shape:
        dim = b"\x08" + _put_varint(int(d))
        out += b"\x12" + _put_varint(len(dim)) + dim
    return out


def _parse_entry(b):
    """BundleEntryProto: dtype=1, shape=2, shard_id=3, offset=4, size=5, crc32c=6 (fixed32), slices=7."""
    e = dict(dtype=0, shape=[], shard_id=0, offset=0, size=0, crc32c=0, slices=0)
    for f, _, v in _pb_fields(b):
        if f == 1:
            e["dtype"] = v
        elif f == 2:
            e["shape"] = _parse_shape(v)
        elif f == 3:
            e["shard_id"] = v
        elif f == 4:
            e["offset"] = v
        elif f == 5:
            e["size"] = v
        elif f == 6:
            e["crc32c"] = v
        elif f == 7:
            e["slices"] += 1
    return e


def _encode_entry(dtype, shape, shard_id, offset, size, crc_masked):
    sh = _encode_shape(shape)
    out = b"\x08" + _put_varint(dtype) + b"\x12" + _put_varint(len(sh)) + sh
    if shard_id:
        out += b"\x18" + _put_varint(shard_id)
    if offset:
        out += b"\x20" + _put_varint(offset)
    out += b"\x28" + _put_varint(size) + b"\x35" + struct.pack("<I", crc_masked)
    return out


# ------------------------------------------------------------------------------------------ LevelDB table
def _read_block(buf, offset, size, verify=True):
    contents = buf[offset:offset + size]
    ctype = buf[offset + size]
    if ctype != 0:
        raise ValueError("compressed table block (type %d): TensorFlow writes checkpoint indices uncompressed" % ctype)
    if verify:
        want = _unmask(struct.unpack_from("<I", buf, offset + size + 1)[0])
        got = crc32c(bytes(buf[offset:offset + size + 1]))
        if want != got:
            raise ValueError("table block checksum mismatch at offset %d" % offset)
    return contents


def _block_entries(block):
    n_restarts = struct.unpack_from("<I", block, len(block) - 4)[0]
    end = len(block) - 4 - 4 * n_restarts
    p, key, out = 0, b"", []
    while p < end:
        shared, p = _get_varint(block, p)
        non_shared, p = _get_varint(block, p)
        vlen, p = _get_varint(block, p)
        key = key[:shared] + bytes(block[p:p + non_shared])
        p += non_shared
        out.append((key, bytes(block[p:p + vlen])))
        p += vlen
    return out


def _table_items(buf, verify=True):
    if len(buf) < 48 or struct.unpack_from("<Q", buf, len(buf) - 8)[0] != _TABLE_MAGIC:
        raise ValueError("not a LevelDB-format table (bad magic)")
    p = len(buf) - 48
    _, p = _get_varint(buf, p)          # metaindex handle
    _, p = _get_varint(buf, p)
    ioff, p = _get_varint(buf, p)
    isize, p = _get_varint(buf, p)
    items = []
    for _, handle in _block_entries(_read_block(buf, ioff, isize, verify)):
        off, q = _get_varint(handle, 0)
        size, _ = _get_varint(handle, q)
        items.extend(_block_entries(_read_block(buf, off, size, verify)))
    return items


def _build_block(items, restart_interval=16):
    out = bytearray()
    restarts = []
    prev = b""
    for i, (k, v) in enumerate(items):
        shared = 0
        if i % restart_interval == 0:
            restarts.append(len(out))
        else:
            m = min(len(prev), len(k))
            while shared < m and prev[shared] == k[shared]:
                shared += 1
        out += _put_varint(shared) + _put_varint(len(k) - shared) + _put_varint(len(v)) + k[shared:] + v
        prev = k
    if not restarts:
        restarts = [0]
    for r in restarts:
        out += struct.pack("<I", r)
    out += struct.pack("<I", len(restarts))
    return bytes(out)


def _build_table(items, block_bytes=4096):
    """Sorted (key, value) pairs -> table file bytes (uncompressed blocks, one index block)."""
    out = bytearray()

    def emit(block):
        off = len(out)
        out.extend(block)
        out.append(0)                                           # kNoCompression
        out.extend(struct.pack("<I", _mask(crc32c(block + b"\x00"))))
        return _put_varint(off) + _put_varint(len(block))

    index, cur, cur_bytes = [], [], 0
    for k, v in items:
        cur.append((k, v))
        cur_bytes += len(k) + len(v) + 3
        if cur_bytes >= block_bytes:
            index.append((cur[-1][0], emit(_build_block(cur))))   # separator = last key of the block
            cur, cur_bytes = [], 0
    if cur:
        index.append((cur[-1][0], emit(_build_block(cur))))
    meta = emit(_build_block([]))
    idx = emit(_build_block(index, restart_interval=1))
    footer = meta + idx
    footer += b"\x00" * (40 - len(footer)) + struct.pack("<Q", _TABLE_MAGIC)
    out.extend(footer)
    return bytes(out)


# ------------------------------------------------------------------------------------------ checkpoints
def read_checkpoint_index(prefix, verify=True):
    """OrderedDict name -> dict(dtype, shape, shard_id, offset, size, crc32c) of `<prefix>.index`, plus
    the header under the key '' (num_shards, version)."""
    path = prefix if prefix.endswith(".index") else prefix + ".index"
    buf = open(path, "rb").read()
    out = OrderedDict()
    for k, v in _table_items(buf, verify):
        if k == b"":
            h = dict(num_shards=1, endianness=0)
            for f, _, val in _pb_fields(v):
                if f == 1:
                    h["num_shards"] = val
                elif f == 2:
                    h["endianness"] = val
            out[""] = h
        else:
            out[k.decode()] = _parse_entry(v)
    return out


def read_checkpoint(prefix, names=None, verify=True):
    """name -> array from a TF-V2 checkpoint (`<prefix>.index` + `<prefix>.data-SSSSS-of-NNNNN`)."""
    idx = read_checkpoint_index(prefix, verify)
    hdr = idx.pop("", dict(num_shards=1))
    if hdr.get("endianness", 0) != 0:
        raise ValueError("big-endian checkpoint")
    shards = {}
    out = OrderedDict()
    for name, e in idx.items():
        if names is not None and name not in names:
            continue
        if e["slices"]:
            raise ValueError("%s: partitioned (sliced) variables are not supported" % name)
        if e["dtype"] not in _NP:
            continue
        sid = e["shard_id"]
        if sid not in shards:
            shards[sid] = np.memmap("%s.data-%05d-of-%05d" % (prefix, sid, hdr["num_shards"]), dtype=np.uint8, mode="r")
        raw = np.asarray(shards[sid][e["offset"]:e["offset"] + e["size"]])
        if verify and _unmask(e["crc32c"]) != crc32c(raw):
            raise ValueError("%s: tensor checksum mismatch" % name)
        out[name] = raw.view(_NP[e["dtype"]]).reshape(e["shape"]).copy()
    return out


def write_checkpoint(prefix, tensors):
    """Write {name: array} as a one-shard TF-V2 checkpoint bundle readable by tf.train.Saver.restore."""
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    items = [(b"", b"\x08\x01\x1a\x02\x08\x01")]         # BundleHeaderProto: num_shards = 1, version { producer: 1 }
    offset = 0
    with open(prefix + ".data-00000-of-00001.tmp", "wb") as f:
        for name in sorted(tensors):
            a = np.asarray(tensors[name])
            a = a if a.flags.c_contiguous else a.copy()         # (ascontiguousarray would turn a 0-d scalar into 1-d)
            if a.dtype not in _TF:
                raise ValueError("%s: dtype %s cannot be stored" % (name, a.dtype))
            raw = a.reshape(-1).view(np.uint8)
            f.write(raw.tobytes())
            items.append((name.encode(), _encode_entry(_TF[a.dtype], a.shape, 0, offset, raw.size, _mask(crc32c(raw)))))
            offset += raw.size
    os.replace(prefix + ".data-00000-of-00001.tmp", prefix + ".data-00000-of-00001")
    with open(prefix + ".index.tmp", "wb") as f:
        f.write(_build_table(items))
    os.replace(prefix + ".index.tmp", prefix + ".index")
    return prefix


# ------------------------------------------------------------------------------------------ frozen graphs
def _tensor_to_numpy(t):
    shape = [d.size for d in t.tensor_shape.dim]
    if t.dtype not in _NP:
        return None
    dt = _NP[t.dtype]
    if t.tensor_content:
        return np.frombuffer(t.tensor_content, dtype=dt).reshape(shape).copy()
    for field in ("float_val", "int_val", "int64_val", "bool_val", "double_val"):
        v = list(getattr(t, field))
        if v:
            a = np.array(v, dtype=dt)
            n = int(np.prod(shape)) if shape else 1
            if a.size == 1 and n > 1:
                a = np.full(n, a[0], dtype=dt)
            return a.reshape(shape)
    return np.zeros(shape, dtype=dt)


def read_frozen_graph(path):
    """{variable name: float32 array} of a frozen GraphDef: `convert_variables_to_constants` turns
    every variable into a Const node that keeps the variable's name (`conv2d_7/kernel`,
    `batch_normalization_3/moving_mean`, `dense/bias`, ...)."""
    from tensorboard.compat.proto import graph_pb2
    g = graph_pb2.GraphDef()
    g.ParseFromString(open(path, "rb").read())
    out = OrderedDict()
    for n in g.node:
        if n.op == "Const" and "/" in n.name:
            a = _tensor_to_numpy(n.attr["value"].tensor)
            if a is not None and a.dtype == np.float32:
                out[n.name] = a
    return out


def write_frozen_graph(path, tensors):
    """The inverse, for weights only: one float32 Const node per variable (enough for
    read_frozen_graph / ResNet.load_frozen_graph and for tf.import_graph_def to resolve the names)."""
    from tensorboard.compat.proto import graph_pb2, tensor_pb2, tensor_shape_pb2
    g = graph_pb2.GraphDef()
    for name, a in tensors.items():
        a = np.ascontiguousarray(a, dtype=np.float32)
        n = g.node.add()
        n.name = name
        n.op = "Const"
        n.attr["dtype"].type = DT_FLOAT
        shape = tensor_shape_pb2.TensorShapeProto(dim=[tensor_shape_pb2.TensorShapeProto.Dim(size=d) for d in a.shape])
        n.attr["value"].tensor.CopyFrom(tensor_pb2.TensorProto(dtype=DT_FLOAT, tensor_shape=shape,
                                                                tensor_content=a.tobytes()))
    with open(path, "wb") as f:
        f.write(g.SerializeToString())
    return path
