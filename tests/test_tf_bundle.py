"""TensorFlow tensor-bundle checkpoints read without TensorFlow.  No real checkpoint exists
in this environment (parity unpinned, see the module header): the reader is checked against the module's own writer,
against hand-assembled bytes of the published layout, and for its error behaviour."""
import os
import struct

import numpy as np
import pytest

from textsummarization_on_flink_b200 import tf_bundle as tb, weights_io, synth


def test_crc32c_known_answers():
    assert tb.crc32c(b"123456789") == 0xE3069283                       # RFC 3720 check value
    assert tb.crc32c(b"\x00" * 32) == 0x8A9136AA
    assert tb.mask_crc(0) == 0xA282EAD8


def test_bundle_round_trip_many_blocks(tmp_path):
    rng = np.random.default_rng(0)
    t = {"seq2seq/embedding/embedding": rng.standard_normal((50, 8)).astype(np.float32),
         "global_step": np.array(237, np.int64), "seq2seq/output_projection/v/Adagrad": np.ones(5, np.float32)}
    for i in range(400):                                                 # several data blocks, shared key prefixes
        t["seq2seq/decoder/attention_decoder/var_%03d" % i] = np.full((2, 3), i, np.float32)
    prefix = str(tmp_path / "model.ckpt-237")
    tb.write_bundle(prefix, t)
    got = tb.read_bundle(prefix, verify_data=True)
    assert set(got) == set(t)
    for k in t:
        assert got[k].dtype == np.asarray(t[k]).dtype and got[k].shape == np.asarray(t[k]).shape
        assert np.array_equal(got[k], t[k])


def test_hand_assembled_index_single_block(tmp_path):
    """One data block written byte by byte from the format description, independent of write_bundle."""
    data = np.arange(6, dtype="<f4").tobytes()
    entry = (b"\x08\x01" + b"\x12\x08" + b"\x12\x02\x08\x02" + b"\x12\x02\x08\x03" + b"\x28\x18" +
             b"\x35" + struct.pack("<I", tb.mask_crc(tb.crc32c(data))))
    header = b"\x08\x01\x10\x00"
    block = (b"\x00\x00" + bytes([len(header)]) + header +                     # key "" (shared 0, unshared 0)
             b"\x00\x01" + bytes([len(entry)]) + b"w" + entry +                # key "w"
             struct.pack("<II", 0, 1))                                           # one restart at 0
    def trailer(b):
        return b"\x00" + struct.pack("<I", tb.mask_crc(tb.crc32c(b + b"\x00")))
    meta = struct.pack("<II", 0, 1)
    f = block + trailer(block)
    meta_off = len(f)
    f += meta + trailer(meta)
    handle = bytes([0, len(block)])
    iblock = b"\x00\x01" + bytes([len(handle)]) + b"w" + handle + struct.pack("<II", 0, 1)
    idx_off = len(f)
    f += iblock + trailer(iblock)
    foot = bytes([meta_off, len(meta), idx_off, len(iblock)])
    f += foot + b"\x00" * (40 - len(foot)) + struct.pack("<Q", tb.TABLE_MAGIC)
    prefix = str(tmp_path / "m")
    open(prefix + ".index", "wb").write(f)
    open(prefix + ".data-00000-of-00001", "wb").write(data)
    got = tb.read_bundle(prefix, verify_data=True)
    assert list(got) == ["w"] and got["w"].shape == (2, 3)
    assert np.array_equal(got["w"], np.arange(6, dtype=np.float32).reshape(2, 3))


def test_corruption_is_detected(tmp_path):
    prefix = str(tmp_path / "m")
    tb.write_bundle(prefix, {"a": np.ones((4, 4), np.float32)})
    raw = bytearray(open(prefix + ".index", "rb").read())
    raw[3] ^= 0x40
    open(prefix + ".index", "wb").write(bytes(raw))
    with pytest.raises(ValueError, match="checksum"):
        tb.read_bundle(prefix)
    open(prefix + ".index", "wb").write(b"not a table")
    with pytest.raises(ValueError, match="magic"):
        tb.read_bundle(prefix)
    tb.write_bundle(prefix, {"a": np.ones((4, 4), np.float32)})
    d = bytearray(open(prefix + ".data-00000-of-00001", "rb").read())
    d[5] ^= 1
    open(prefix + ".data-00000-of-00001", "wb").write(bytes(d))
    with pytest.raises(ValueError, match="data checksum"):
        tb.read_bundle(prefix, verify_data=True)


def test_convert_checkpoint_to_pgenw(tmp_path, small_cfg):
    """util.py:29-41: the newest checkpoint named by the ``checkpoint`` file; optimiser slots are dropped; the result
    is the archive the library loads (same names, same values)."""
    W = synth.make_weights("W1", vsize=small_cfg["V"], hidden=small_cfg["H"], emb=small_cfg["E"], seed=3)
    t = dict(W)
    t["global_step"] = np.array(1000, np.int64)
    for n in list(W)[:3]:
        t[n + "/Adagrad"] = np.full_like(W[n], 0.1)
    train = tmp_path / "train"
    train.mkdir()
    tb.write_bundle(str(train / "model.ckpt-1000"), t)
    (train / "checkpoint").write_text('model_checkpoint_path: "model.ckpt-1000"\nall_model_checkpoint_paths: "model.ckpt-1000"\n')
    out = str(train / "model.pgenw")
    names = tb.convert_checkpoint(str(train), out)
    assert names == sorted(W)
    back = weights_io.load_weights(out)
    for n in W:
        assert np.array_equal(back[n], W[n])
