"""The stage-level contract of TFModel (params, defaults, JSON round trip, schema, errors), mirrored from
src/test/java/.../TensorFlowTest.java -- the part of it that needs no GPU -- plus the GPU run of
testModelInference's 8 rows."""
import json

import pytest

import pgen_oracle as O
from conftest import small_flags
from textsummarization_on_flink_b200 import pipeline, synth


def create_model(hyper):
    """TensorFlowTest.createModel(), :219-234."""
    return (pipeline.TFModel()
            .setZookeeperConnStr("127.0.0.1:2181").setWorkerNum(1).setPsNum(0)
            .setInferenceScripts(["run_summarization.py"]).setInferenceMapFunc("main_on_flink")
            .setInferenceHyperParamsKey("TF_Hyperparameter").setInferenceHyperParams(hyper)
            .setInferenceEnvPath(None)
            .setInferenceSelectedCols(["uuid", "article", "reference"])
            .setInferenceOutputCols(["uuid", "article", "summary", "reference"])
            .setInferenceOutputTypes([pipeline.STRING] * 4))


def test_param_defaults_and_fluent_setters():
    m = pipeline.TFModel()
    assert m.getZookeeperConnStr() == "127.0.0.1:2181" and m.getWorkerNum() == 1 and m.getPsNum() == 0
    assert m.getInferenceHyperParams() == [] and m.getInferenceEnvPath() is None
    with pytest.raises(RuntimeError):
        m.getInferenceScripts()                      # required, no default (HasInferencePythonConfig.java:17-20)
    assert m.setWorkerNum(8) is m and m.getWorkerNum() == 8


def test_json_round_trip():
    hyper = ["run_summarization.py", "--mode=decode", "--coverage=1", "--single_pass=1", "--inference=1"]
    m = create_model(hyper)
    s = m.toJson()
    m2 = pipeline.TFModel().loadJson(s)              # TensorFlowTest.java:152-163
    assert m2.getParams() == m.getParams() and json.loads(s)["inference_hyper_params"] == hyper
    assert m2.hyperparams_string() == " ".join(hyper)
    with pytest.raises(RuntimeError):
        pipeline.TFModel().loadJson('{"bogus": 1}')


def test_schema_and_flag_errors_surface_as_runtime_error():
    rows = [{"uuid": "u", "article": "a", "reference": "r"}]
    with pytest.raises(RuntimeError):                # selected cols must carry the three features
        list(create_model(["p", "--mode=decode"]).setInferenceSelectedCols(["uuid", "article"]).transform(rows))
    with pytest.raises(RuntimeError):
        list(create_model(["p", "--mode=decode"]).setInferenceOutputTypes(["STRING"] * 3).transform(rows))
    with pytest.raises(RuntimeError) as e:           # mode=train is not served; RuntimeException in the reference
        list(create_model(["p", "--mode=train"]).transform(rows))
    assert "decode" in str(e.value)
    with pytest.raises(RuntimeError):
        list(create_model(["p", "--mode=decode", "--vocab_path=/nonexistent"]).transform(rows))


@pytest.mark.gpu
def test_model_inference_rows(small_cfg, small_model_dir):
    """TensorFlowTest.testModelInference (:93-107) with createArticleData() (:204-214): 8 rows
    ("uuid-i", "article i.", "", "reference i."), plus testPipeline's extra column being dropped by the select."""
    flags = small_flags(small_model_dir, small_cfg).split(" ")
    rows = [{"uuid": "uuid-%d" % i, "article": "article %d." % i, "summary": "", "reference": "reference %d." % i}
            for i in range(8)]
    out = list(create_model(flags).transform(rows))
    assert [r["uuid"] for r in out] == [b"uuid-%d" % i for i in range(8)]
    hps = O.parse_hyperparams(" ".join(flags))
    vocab = O.Vocab(small_model_dir["vocab_path"], small_cfg["V"])
    for i, r in enumerate(out):
        assert list(r) == ["uuid", "article", "summary", "reference"]
        ref = O.decode_row(small_cfg["W"], vocab, hps, b"uuid-%d" % i, b"article %d." % i, b"reference %d." % i, mode="dedup")
        assert (r["uuid"], r["article"], r["summary"], r["reference"]) == ref
    # worker_num = 2 on one device: same rows, same order
    out2 = list(create_model(flags).setWorkerNum(2).transform(rows, devices=[0]))
    assert out2 == out
    # output schema is honoured
    out3 = list(create_model(flags).setInferenceOutputCols(["uuid", "summary"]).setInferenceOutputTypes([pipeline.STRING] * 2).transform(rows))
    assert [list(r) for r in out3] == [["uuid", "summary"]] * 8 and [r["summary"] for r in out3] == [r["summary"] for r in out]


@pytest.mark.gpu
def test_two_subtasks_on_two_devices_in_one_process(model_dir):
    """worker_num = 2 with one GPU per subtask inside ONE process (two handles on two devices): full-size model, so
    every per-device kernel attribute and descriptor is exercised on both devices.  Needs two GPUs."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    flags = synth.hyperparams_string(model_dir, extra="--articles_per_batch=4").split(" ")
    arts = synth.make_article_ids(8, seed=17)
    rows = [{"uuid": "u%d" % i, "article": synth.ids_to_article(e, 50000, salt=i), "reference": "r"} for i, (_, e, _) in enumerate(arts)]
    one = list(create_model(flags).transform(rows, devices=[0]))
    two = list(create_model(flags).setWorkerNum(2).transform(rows, devices=[0, 1]))
    assert two == one
