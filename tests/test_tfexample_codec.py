"""tf.Example wire codec of libpgen_cuda.so against google.protobuf (an independent implementation of the wire
format; TensorFlow's example.proto is restated below from its published definition).  No GPU needed."""
import pytest
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

from textsummarization_on_flink_b200 import native


def _example_classes():
    fd = descriptor_pb2.FileDescriptorProto(name="pgen_test_example.proto", package="pgentest", syntax="proto3")
    T = descriptor_pb2.FieldDescriptorProto

    def msg(name, fields, nested=()):
        m = fd.message_type.add(name=name)
        for fname, num, typ, label, tname in fields:
            f = m.field.add(name=fname, number=num, type=typ, label=label)
            if tname:
                f.type_name = tname
        return m

    msg("BytesList", [("value", 1, T.TYPE_BYTES, T.LABEL_REPEATED, None)])
    msg("FloatList", [("value", 1, T.TYPE_FLOAT, T.LABEL_REPEATED, None)])
    msg("Int64List", [("value", 1, T.TYPE_INT64, T.LABEL_REPEATED, None)])
    feat = msg("Feature", [("bytes_list", 1, T.TYPE_MESSAGE, T.LABEL_OPTIONAL, ".pgentest.BytesList"),
                           ("float_list", 2, T.TYPE_MESSAGE, T.LABEL_OPTIONAL, ".pgentest.FloatList"),
                           ("int64_list", 3, T.TYPE_MESSAGE, T.LABEL_OPTIONAL, ".pgentest.Int64List")])
    feat.oneof_decl.add(name="kind")
    for f in feat.field:
        f.oneof_index = 0
    feats = fd.message_type.add(name="Features")
    entry = feats.nested_type.add(name="FeatureEntry")
    entry.options.map_entry = True
    entry.field.add(name="key", number=1, type=T.TYPE_STRING, label=T.LABEL_OPTIONAL)
    entry.field.add(name="value", number=2, type=T.TYPE_MESSAGE, label=T.LABEL_OPTIONAL, type_name=".pgentest.Feature")
    feats.field.add(name="feature", number=1, type=T.TYPE_MESSAGE, label=T.LABEL_REPEATED,
                    type_name=".pgentest.Features.FeatureEntry")
    msg("Example", [("features", 1, T.TYPE_MESSAGE, T.LABEL_OPTIONAL, ".pgentest.Features")])
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("pgentest.Example"))


Example = _example_classes()


def pb_example(**feats):
    ex = Example()
    for k, v in feats.items():
        if isinstance(v, bytes):
            ex.features.feature[k].bytes_list.value.append(v)
        elif isinstance(v, list) and v and isinstance(v[0], bytes):
            ex.features.feature[k].bytes_list.value.extend(v)
        elif isinstance(v, float):
            ex.features.feature[k].float_list.value.append(v)
        else:
            ex.features.feature[k].int64_list.value.append(v)
    return ex.SerializeToString()


CASES = [(b"uuid-0", b"article 0.", b"reference 0."), (b"", b"", b""),
         (b"u" * 300, bytes(range(256)) * 40, b"\x00\x01 ref \xff"), (b"\xe4\xb8\xad", "文章 。".encode(), b"r")]


@pytest.mark.parametrize("uuid,article,reference", CASES)
def test_parse_matches_protobuf_encoder(uuid, article, reference):
    buf = pb_example(uuid=uuid, article=article, reference=reference)
    assert native.example_parse(buf) == (uuid, article, b"", reference)
    # extra features of other kinds and an extra bytes feature are ignored; summary is picked up when present
    buf = pb_example(uuid=uuid, weight=0.5, article=article, n=7, reference=reference, summary=b"s .", extra=b"x")
    assert native.example_parse(buf) == (uuid, article, b"s .", reference)


@pytest.mark.parametrize("uuid,article,reference", CASES)
def test_write_is_decoded_by_protobuf(uuid, article, reference):
    summary = b"a b .  c ."
    buf = native.example_write(uuid, article, summary, reference)
    ex = Example.FromString(buf)
    got = {k: list(v.bytes_list.value) for k, v in ex.features.feature.items()}
    assert got == {"uuid": [uuid], "article": [article], "summary": [summary], "reference": [reference]}
    assert native.example_parse(buf) == (uuid, article, summary, reference)          # round trip
    # protobuf re-serialisation of the same message has the same size (map order may differ)
    assert len(ex.SerializeToString()) == len(buf)


def test_parse_errors_like_fixed_len_feature():
    with pytest.raises(native.PgenError) as e:                     # missing feature (batcher.py:554-557 would raise)
        native.example_parse(pb_example(uuid=b"u", article=b"a"))
    assert "reference" in str(e.value)
    with pytest.raises(native.PgenError):                          # two values where FixedLenFeature([]) wants one
        native.example_parse(pb_example(uuid=b"u", article=[b"a", b"b"], reference=b"r"))
    with pytest.raises(native.PgenError):                          # wrong kind
        native.example_parse(pb_example(uuid=b"u", article=3, reference=b"r"))
    good = pb_example(uuid=b"u", article=b"a" * 50, reference=b"r")
    for cut in (1, 5, len(good) // 2, len(good) - 1):              # truncated buffers never crash
        with pytest.raises(native.PgenError):
            native.example_parse(good[:cut])
    with pytest.raises(native.PgenError):
        native.example_parse(b"\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff")


@pytest.mark.gpu
def test_submit_example_rows(small_cfg, small_model_dir):
    import pgen_oracle as O
    from conftest import small_flags
    flags = small_flags(small_model_dir, small_cfg, extra="--articles_per_batch=4")
    hps = O.parse_hyperparams(flags)
    vocab = O.Vocab(small_model_dir["vocab_path"], small_cfg["V"])
    rows = [(b"uuid-%d" % i, b"article %d. w7 w9 . w11" % i, b"reference %d." % i) for i in range(6)]
    with native.Decoder(flags) as dec:
        for u, a, r in rows:
            dec.submit_example(pb_example(uuid=u, article=a, reference=r))
        out = dec.poll(flush=True)
        with pytest.raises(native.PgenError):
            dec.submit_example(pb_example(uuid=b"u", article=b"a"))
    for (u, a, r), got in zip(rows, out):
        ref = O.decode_row(small_cfg["W"], vocab, hps, u, a, r, mode="dedup")
        assert got == ref
        ex = Example.FromString(native.example_write(*got))       # what goes back through the output queue
        assert ex.features.feature["summary"].bytes_list.value[0] == ref[2]
