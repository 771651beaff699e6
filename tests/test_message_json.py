"""the demo application's JSON records either side of the stage (Message.java,
MessageDeserializationSchema.java, MessageSerializationSchema.java).  No JVM here: checked against the classes' source
and Jackson's documented defaults."""
import json

import pytest

from conftest import small_flags
from textsummarization_on_flink_b200 import message, pipeline


def test_serialize_field_order_escapes_and_null():
    out = message.serialize(["u-1", 'a "quoted" \\ line\nnext\ttab \x01 é 中', None, "r/ef"])
    assert out == ('{"uuid":"u-1","article":"a \\"quoted\\" \\\\ line\\nnext\\ttab \\u0001 é 中","summary":null,'
                   '"reference":"r/ef"}').encode("utf-8")
    assert json.loads(out) == {"uuid": "u-1", "article": 'a "quoted" \\ line\nnext\ttab \x01 é 中', "summary": None,
                               "reference": "r/ef"}
    assert message.serialize([b"u", b"a", b"s", b"r"]) == b'{"uuid":"u","article":"a","summary":"s","reference":"r"}'
    assert message.serialize(["only", "three", "fields"]) == b""           # logged and dropped in the reference
    assert message.serialize([1, "a", "s", "r"]) == b""


def test_deserialize_defaults_and_errors():
    de = message.MessageDeserializer(max_count=2)
    assert de.deserialize(b'{"uuid":"u","article":"a","summary":"","reference":"r"}') == ("u", "a", "", "r")
    assert de.deserialize(b'{"reference":"r","uuid":"u"}') == ("u", None, None, "r")      # missing -> null
    assert not de.is_end_of_stream()
    assert de.deserialize(b'{"uuid":7,"article":true,"summary":null,"reference":"r"}') == ("7", "true", None, "r")
    assert de.is_end_of_stream()                                            # counter (3) > maxCount (2)
    with pytest.raises(ValueError, match="Unrecognized field"):
        de.deserialize(b'{"uuid":"u","title":"x"}')
    with pytest.raises(ValueError):
        de.deserialize(b'{"uuid":{"a":1}}')
    with pytest.raises(ValueError):
        de.deserialize(b'["u","a","s","r"]')
    with pytest.raises(ValueError):
        de.deserialize(b'{"uuid":')


def test_round_trip():
    rows = [("u%d" % i, "article %d ." % i, "", "ref é %d" % i) for i in range(5)]
    de = message.MessageDeserializer(100)
    assert [de.deserialize(message.serialize(r)) for r in rows] == rows


@pytest.mark.gpu
def test_json_records_through_the_stage(small_cfg, small_model_dir):
    """App.startInference without Kafka: JSON in, JSON out, same summaries as the row API, source ends after
    max_count records."""
    flags = small_flags(small_model_dir, small_cfg).split(" ")

    def model():
        return (pipeline.TFModel().setInferenceScripts(["run_summarization.py"]).setInferenceMapFunc("main_on_flink")
                .setInferenceHyperParamsKey("TF_Hyperparameter").setInferenceHyperParams(flags)
                .setInferenceSelectedCols(["uuid", "article", "reference"])
                .setInferenceOutputCols(["uuid", "article", "summary", "reference"])
                .setInferenceOutputTypes([pipeline.STRING] * 4))
    rows = [{"uuid": "uuid-%d" % i, "article": "article %d." % i, "reference": "reference %d." % i} for i in range(6)]
    want = list(model().transform(rows))
    recs = [message.serialize([r["uuid"], r["article"], "", r["reference"]]) for r in rows]
    got = [json.loads(b) for b in message.transform_messages(model(), recs, max_count=4)]
    assert len(got) == 4                                                    # records 5 and 6 are past the limit
    for g, w in zip(got, want):
        assert g == {k: w[k].decode() for k in message.FIELDS}
