"""Flat weights archive (``.pgenw``) read by libpgen_cuda.so and written from NumPy.

The reference restores every graph variable from the newest TF checkpoint under
``<log_root>/<exp_name>/train`` (decode.py:58-65, util.py:29-41).  TensorFlow is not
available, so the library reads a flat archive instead, keyed by the *same* TF variable
names so that a real checkpoint can be converted 1:1.

Layout (little endian)::

    char[8]  magic  "PGENW001"
    u32      n_tensors
    n x { u16 name_len; char name[name_len]; u8 dtype (0 = float32); u8 ndim;
          u32 dims[ndim]; u64 offset; u64 nbytes }
    ...      tensor data, each tensor 64-byte aligned, C order
"""
from __future__ import annotations

import struct
from typing import Dict

import numpy as np

MAGIC = b"PGENW001"
DEFAULT_NAME = "model.pgenw"


def save_weights(path: str, tensors: Dict[str, np.ndarray]) -> None:
    names = sorted(tensors)
    head = 8 + 4
    for n in names:
        head += 2 + len(n.encode()) + 1 + 1 + 4 * tensors[n].ndim + 8 + 8
    off = (head + 63) // 64 * 64
    entries = []
    for n in names:
        a = np.ascontiguousarray(tensors[n], dtype=np.float32)
        entries.append((n, a, off))
        off = (off + a.nbytes + 63) // 64 * 64
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<I", len(entries)))
        for n, a, o in entries:
            nb = n.encode()
            f.write(struct.pack("<H", len(nb)))
            f.write(nb)
            f.write(struct.pack("<BB", 0, a.ndim))
            f.write(struct.pack("<%dI" % a.ndim, *a.shape))
            f.write(struct.pack("<QQ", o, a.nbytes))
        for n, a, o in entries:
            f.seek(o)
            f.write(a.tobytes())
        f.truncate(off)


def load_weights(path: str) -> Dict[str, np.ndarray]:
    with open(path, "rb") as f:
        buf = f.read()
    if buf[:8] != MAGIC:
        raise ValueError("%s: not a .pgenw archive" % path)
    (n,) = struct.unpack_from("<I", buf, 8)
    p = 12
    out = {}
    for _ in range(n):
        (ln,) = struct.unpack_from("<H", buf, p)
        p += 2
        name = buf[p:p + ln].decode()
        p += ln
        dtype, ndim = struct.unpack_from("<BB", buf, p)
        p += 2
        dims = struct.unpack_from("<%dI" % ndim, buf, p)
        p += 4 * ndim
        off, nbytes = struct.unpack_from("<QQ", buf, p)
        p += 16
        if dtype != 0:
            raise ValueError("unsupported dtype code %d" % dtype)
        out[name] = np.frombuffer(buf, dtype=np.float32, count=nbytes // 4, offset=off).reshape(dims).copy()
    return out
