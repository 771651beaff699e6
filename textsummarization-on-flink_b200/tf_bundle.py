"""TensorFlow checkpoint (tensor bundle, "V2" format) reader/writer without TensorFlow -- SURVEY.md section 8f row 2.

The reference restores the newest checkpoint under ``<log_root>/<exp_name>/train`` (util.py:29-41, decode.py:58-65);
the published ``pretrained_model_tf1.2.1`` (log/download_model.sh:14-28) is such a bundle.  This module turns one into
the flat ``.pgenw`` archive the library reads, keyed by the same variable names.

PARITY UNPINNED: there is no TensorFlow and no checkpoint file in this environment, so the format below is restated
from its published description (tensorflow/core/util/tensor_bundle, tensorflow/core/lib/io/table = the LevelDB table
format) and tested only against this module's own writer.

``<prefix>.index``  LevelDB-style table, keys sorted:
    data block   = entries {varint shared, varint unshared, varint value_len, key suffix, value} ... then
                   u32 restart offsets[], u32 n_restarts; every block is followed by a 5-byte trailer
                   {u8 compression (0 = none), u32 masked crc32c(block + compression byte)}
    index block  = same layout, value = BlockHandle {varint offset, varint size} of a data block
    footer (48B) = metaindex BlockHandle, index BlockHandle, zero padding to 40 bytes, u64 magic 0xdb4775248b80fb57
    key ""       -> BundleHeaderProto {1: num_shards, 2: endianness, 3: version}
    key <name>   -> BundleEntryProto {1: dtype, 2: TensorShapeProto{2: Dim{1: size}}, 3: shard_id, 4: offset, 5: size,
                                      6: fixed32 masked crc32c of the tensor bytes}
``<prefix>.data-NNNNN-of-MMMMM``  raw little-endian tensor bytes.
"""
from __future__ import annotations

import os
import re
import struct
from typing import Dict, Iterator, List, Tuple

import numpy as np

TABLE_MAGIC = 0xDB4775248B80FB57
DT_FLOAT, DT_DOUBLE, DT_INT32, DT_INT64 = 1, 2, 3, 9
_DTYPES = {DT_FLOAT: np.dtype("<f4"), DT_DOUBLE: np.dtype("<f8"), DT_INT32: np.dtype("<i4"), DT_INT64: np.dtype("<i8")}
_DT_OF = {v: k for k, v in _DTYPES.items()}

# ---------------------------------------------------------------- crc32c (Castagnoli), masked as LevelDB / TF do
_POLY = 0x82F63B78
_TABLE = np.zeros(256, np.uint32)
for _i in range(256):
    _c = _i
    for _ in range(8):
        _c = (_c >> 1) ^ _POLY if _c & 1 else _c >> 1
    _TABLE[_i] = _c
_TABLE_L = [int(x) for x in _TABLE]


def crc32c(data: bytes, crc: int = 0) -> int:
    c = crc ^ 0xFFFFFFFF
    t = _TABLE_L
    for b in data:
        c = t[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def mask_crc(c: int) -> int:
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


# ---------------------------------------------------------------- varints and the tiny protobuf subset
def _put_varint(v: int) -> bytes:
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _get_varint(buf: bytes, p: int) -> Tuple[int, int]:
    v, shift = 0, 0
    while True:
        if p >= len(buf):
            raise ValueError("truncated varint")
        b = buf[p]
        p += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, p
        shift += 7
        if shift > 63:
            raise ValueError("varint too long")


def _fields(buf: bytes) -> Iterator[Tuple[int, int, object]]:
    """(field number, wire type, value) of one protobuf message; wire types 0, 1, 2, 5."""
    p = 0
    while p < len(buf):
        tag, p = _get_varint(buf, p)
        fn, wt = tag >> 3, tag & 7
        if wt == 0:
            v, p = _get_varint(buf, p)
        elif wt == 1:
            v = struct.unpack_from("<Q", buf, p)[0]
            p += 8
        elif wt == 2:
            n, p = _get_varint(buf, p)
            if p + n > len(buf):
                raise ValueError("truncated length-delimited field")
            v = buf[p:p + n]
            p += n
        elif wt == 5:
            v = struct.unpack_from("<I", buf, p)[0]
            p += 4
        else:
            raise ValueError("unsupported protobuf wire type %d" % wt)
        yield fn, wt, v


def _signed64(v: int) -> int:
    return v - (1 << 64) if v >= 1 << 63 else v


def _parse_entry(buf: bytes) -> dict:
    e = {"dtype": 0, "shape": [], "shard_id": 0, "offset": 0, "size": 0, "crc32c": None, "sliced": False}
    for fn, wt, v in _fields(buf):
        if fn == 1:
            e["dtype"] = v
        elif fn == 2:
            for f2, _, v2 in _fields(v):
                if f2 == 2:                                 # Dim
                    size = 0
                    for f3, _, v3 in _fields(v2):
                        if f3 == 1:
                            size = _signed64(v3)
                    e["shape"].append(size)
        elif fn == 3:
            e["shard_id"] = v
        elif fn == 4:
            e["offset"] = v
        elif fn == 5:
            e["size"] = v
        elif fn == 6:
            e["crc32c"] = v
        elif fn == 7:
            e["sliced"] = True
    return e


def _entry_proto(dtype: int, shape, shard_id: int, offset: int, size: int, crc: int) -> bytes:
    dims = b"".join(b"\x12" + _put_varint(len(d)) + d for d in (b"\x08" + _put_varint(int(s)) for s in shape))
    out = b"\x08" + _put_varint(dtype) + b"\x12" + _put_varint(len(dims)) + dims
    if shard_id:
        out += b"\x18" + _put_varint(shard_id)
    if offset:
        out += b"\x20" + _put_varint(offset)
    out += b"\x28" + _put_varint(size) + b"\x35" + struct.pack("<I", crc)
    return out


# ---------------------------------------------------------------- the table (.index)
def _read_block(buf: bytes, offset: int, size: int, verify: bool) -> bytes:
    if offset + size + 5 > len(buf):
        raise ValueError("block handle outside the file")
    block, ctype = buf[offset:offset + size], buf[offset + size]
    if verify:
        want = struct.unpack_from("<I", buf, offset + size + 1)[0]
        if mask_crc(crc32c(buf[offset:offset + size + 1])) != want:
            raise ValueError("table block checksum mismatch at offset %d" % offset)
    if ctype != 0:
        raise ValueError("compressed table block (type %d): only uncompressed bundles are supported" % ctype)
    return block


def _block_entries(block: bytes) -> Iterator[Tuple[bytes, bytes]]:
    if len(block) < 4:
        raise ValueError("table block too small")
    n_restarts = struct.unpack_from("<I", block, len(block) - 4)[0]
    end = len(block) - 4 - 4 * n_restarts
    if end < 0:
        raise ValueError("bad restart array")
    p, key = 0, b""
    while p < end:
        shared, p = _get_varint(block, p)
        unshared, p = _get_varint(block, p)
        vlen, p = _get_varint(block, p)
        if shared > len(key) or p + unshared + vlen > end:
            raise ValueError("corrupt table entry")
        key = key[:shared] + block[p:p + unshared]
        p += unshared
        yield key, block[p:p + vlen]
        p += vlen


def read_index(path: str, verify: bool = True) -> Dict[str, dict]:
    """{tensor name: entry dict}; the header entry (key "") is returned under ""."""
    with open(path, "rb") as f:
        buf = f.read()
    if len(buf) < 48 or struct.unpack_from("<Q", buf, len(buf) - 8)[0] != TABLE_MAGIC:
        raise ValueError("%s: not a tensor-bundle index (bad table magic)" % path)
    foot = buf[len(buf) - 48:len(buf) - 8]
    p = 0
    _, p = _get_varint(foot, p)          # metaindex handle: offset, size
    _, p = _get_varint(foot, p)
    ioff, p = _get_varint(foot, p)
    isize, p = _get_varint(foot, p)
    out: Dict[str, dict] = {}
    for _, handle in _block_entries(_read_block(buf, ioff, isize, verify)):
        boff, q = _get_varint(handle, 0)
        bsize, q = _get_varint(handle, q)
        for key, value in _block_entries(_read_block(buf, boff, bsize, verify)):
            if key == b"":
                hdr = {"num_shards": 1, "endianness": 0}
                for fn, _, v in _fields(value):
                    if fn == 1:
                        hdr["num_shards"] = v
                    elif fn == 2:
                        hdr["endianness"] = v
                out[""] = hdr
            else:
                out[key.decode()] = _parse_entry(value)
    if "" not in out:
        raise ValueError("%s: bundle header entry missing" % path)
    if out[""]["endianness"] != 0:
        raise ValueError("big-endian bundles are not supported")
    return out


def _build_block(items: List[Tuple[bytes, bytes]], restart_interval: int = 16) -> bytes:
    out, restarts, prev = bytearray(), [], b""
    for i, (k, v) in enumerate(items):
        shared = 0
        if i % restart_interval == 0:
            restarts.append(len(out))
        else:
            while shared < min(len(prev), len(k)) and prev[shared] == k[shared]:
                shared += 1
        out += _put_varint(shared) + _put_varint(len(k) - shared) + _put_varint(len(v)) + k[shared:] + v
        prev = k
    if not restarts:
        restarts = [0]
    out += struct.pack("<%dI" % len(restarts), *restarts) + struct.pack("<I", len(restarts))
    return bytes(out)


def _emit_block(f, block: bytes) -> Tuple[int, int]:
    off = f.tell()
    f.write(block + b"\x00" + struct.pack("<I", mask_crc(crc32c(block + b"\x00"))))
    return off, len(block)


# ---------------------------------------------------------------- bundles
def read_bundle(prefix: str, verify_data: bool = False) -> Dict[str, np.ndarray]:
    """All whole (un-sliced) tensors of the bundle ``prefix`` (``prefix.index`` + ``prefix.data-*``)."""
    index = read_index(prefix + ".index")
    n_shards = index[""]["num_shards"]
    shards: Dict[int, np.memmap] = {}
    out: Dict[str, np.ndarray] = {}
    for name, e in index.items():
        if name == "" or e["sliced"]:
            continue
        if e["dtype"] not in _DTYPES:
            continue                                        # strings / resources: nothing the model needs
        sid = e["shard_id"]
        if sid not in shards:
            shards[sid] = np.memmap("%s.data-%05d-of-%05d" % (prefix, sid, n_shards), dtype=np.uint8, mode="r")
        raw = shards[sid][e["offset"]:e["offset"] + e["size"]]
        dt = _DTYPES[e["dtype"]]
        count = int(np.prod(e["shape"], dtype=np.int64)) if e["shape"] else 1
        if raw.size != e["size"] or count * dt.itemsize != e["size"]:
            raise ValueError("tensor '%s': %d bytes in the index, shape %s needs %d" % (name, e["size"], e["shape"], count * dt.itemsize))
        if verify_data and e["crc32c"] is not None and mask_crc(crc32c(raw.tobytes())) != e["crc32c"]:
            raise ValueError("tensor '%s': data checksum mismatch" % name)
        out[name] = np.frombuffer(raw.tobytes(), dtype=dt).reshape(e["shape"])
    return out


def write_bundle(prefix: str, tensors: Dict[str, np.ndarray], block_bytes: int = 4096) -> None:
    """Single-shard bundle; used by the tests and to hand weights to TensorFlow-side tools."""
    names = sorted(tensors, key=lambda s: s.encode())
    items: List[Tuple[bytes, bytes]] = [(b"", b"\x08\x01\x10\x00\x1a\x02\x08\x01")]   # 1 shard, little endian, producer 1
    with open(prefix + ".data-00000-of-00001", "wb") as f:
        for n in names:
            a = np.asarray(tensors[n])
            a = np.ascontiguousarray(a).reshape(a.shape)          # ascontiguousarray promotes 0-d to 1-d
            if a.dtype.newbyteorder("<") not in _DT_OF and a.dtype not in _DT_OF:
                raise ValueError("unsupported dtype %s for '%s'" % (a.dtype, n))
            a = a.astype(a.dtype.newbyteorder("<"), copy=False)
            raw = a.tobytes()
            items.append((n.encode(), _entry_proto(_DT_OF[np.dtype(a.dtype.str)], a.shape, 0, f.tell(), len(raw),
                                                   mask_crc(crc32c(raw)))))
            f.write(raw)
    with open(prefix + ".index", "wb") as f:
        handles, cur, size = [], [], 0
        for k, v in items:
            cur.append((k, v))
            size += len(k) + len(v)
            if size >= block_bytes:
                handles.append((cur[-1][0], _emit_block(f, _build_block(cur))))
                cur, size = [], 0
        if cur:
            handles.append((cur[-1][0], _emit_block(f, _build_block(cur))))
        meta = _emit_block(f, _build_block([]))
        idx = _emit_block(f, _build_block([(k, _put_varint(o) + _put_varint(s)) for k, (o, s) in handles], 1))
        foot = _put_varint(meta[0]) + _put_varint(meta[1]) + _put_varint(idx[0]) + _put_varint(idx[1])
        f.write(foot + b"\x00" * (40 - len(foot)) + struct.pack("<Q", TABLE_MAGIC))


def latest_checkpoint(train_dir: str) -> str:
    """tf.train.get_checkpoint_state(train_dir).model_checkpoint_path (util.py:34-36): the ``checkpoint`` text file."""
    with open(os.path.join(train_dir, "checkpoint")) as f:
        m = re.search(r'^model_checkpoint_path:\s*"(.*)"\s*$', f.read(), re.M)
    if not m:
        raise ValueError("no model_checkpoint_path in %s/checkpoint" % train_dir)
    p = m.group(1)
    return p if os.path.isabs(p) else os.path.join(train_dir, p)


def convert_checkpoint(prefix_or_dir: str, out_path: str) -> List[str]:
    """Checkpoint -> .pgenw.  Optimiser slots (``.../Adagrad``) and non-float bookkeeping (``global_step``) are
    dropped, as the reference's decode-mode Saver only needs the model variables (decode.py:58-65)."""
    from . import weights_io
    prefix = latest_checkpoint(prefix_or_dir) if os.path.isdir(prefix_or_dir) else prefix_or_dir
    keep = {n: a.astype(np.float32) for n, a in read_bundle(prefix).items()
            if a.dtype == np.float32 and not n.endswith("/Adagrad") and "/Adagrad_" not in n}
    if not keep:
        raise ValueError("%s: no float32 model variables found" % prefix)
    weights_io.save_weights(out_path, keep)
    return sorted(keep)
