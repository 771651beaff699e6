"""Host logic of libpgen_cuda.so (csrc/capi.cu, csrc/engine.cu) on the CPU: the library's own objects (csrc/build/*.o,
unmodified) linked against a TEST DOUBLE of the CUDA runtime (tests/fake_cuda/fake_cudart.c: "device" memory is zeroed
host memory, copies are memcpy, kernel and graph launches do nothing, any CUDA call can be told to fail).  Nothing is
decoded -- every summary comes back empty -- so this says nothing about kernels or parity; it covers what the GPU tests
cannot or do not: the row state machine (batching, slots, order, flush deadline, the stage API being refused while
rows are in flight) and, above all, the ERROR paths: a failing CUDA call must surface as PGEN_ERR_CUDA and must not
lose rows (MLMapFunction.flatMap's contract: every row that went in comes out)."""
import ctypes as C
import importlib.util
import os
import subprocess
import time

import numpy as np
import pytest

from conftest import small_flags
from textsummarization_on_flink_b200 import native as real_native
from textsummarization_on_flink_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "textsummarization-on-flink_b200", "csrc")


@pytest.fixture(scope="module")
def fake(tmp_path_factory):
    """-> (native module bound to the fake-runtime build, the fake runtime's control functions)"""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "fake_cuda"))
    import build_fake
    return build_fake.build(str(tmp_path_factory.mktemp("fake_cuda")))


def _flags(small_model_dir, small_cfg, extra):
    return small_flags(small_model_dir, small_cfg, extra=extra)


def _articles(n, V):
    return synth.make_articles(n, vsize=V, seed=77, max_len=24, oov_rate=0.1)


def test_rows_come_back_in_order_with_partial_batches_and_slot_rotation(fake, small_cfg, small_model_dir):
    nat, L = fake
    arts = _articles(23, small_cfg["V"])
    with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=4 --slots=2 --flush_ms=-1")) as dec:
        out = []
        for i, a in enumerate(arts):
            dec.submit(b"u%d" % i, a, b"ref %d" % i)
            out += dec.poll()
        assert len(out) >= 12                      # full batches came back without any flush
        out += dec.poll(flush=True)               # 23 = 5 x 4 + 3: the partial batch needs the flush
        st = dec.stats()
    assert [r[0] for r in out] == [b"u%d" % i for i in range(23)]
    assert [r[1] for r in out] == arts and [r[3] for r in out] == [b"ref %d" % i for i in range(23)]
    assert all(r[2] == b"" for r in out)          # the test double decodes nothing
    assert st["rows_in"] == st["rows_out"] == 23 and st["batches"] == 6 and st["kernel_launches"] > 0
    assert L.fake_cuda_launches() > 0 and L.fake_cuda_graph_launches() > 0


def test_flush_deadline_and_no_deadline(fake, small_cfg, small_model_dir):
    nat, _ = fake
    art = _articles(1, small_cfg["V"])[0]
    with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=8 --flush_ms=30")) as dec:
        dec.submit(b"lone", art, b"")
        assert dec.poll() == []                    # deadline not reached
        time.sleep(0.05)
        got = dec.poll() or dec.poll()             # the deadline launches the partial batch; it is finished at once here
        assert [r[0] for r in got] == [b"lone"]
    with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=8 --flush_ms=-1")) as dec:
        dec.submit(b"lone", art, b"")
        time.sleep(0.05)
        assert dec.poll() == []                    # flush_ms=-1: only flush=1 (end of input) decodes a partial batch
        assert [r[0] for r in dec.poll(flush=True)] == [b"lone"]


def test_stage_api_refused_while_rows_in_flight(fake, small_cfg, small_model_dir):
    nat, _ = fake
    arts = _articles(2, small_cfg["V"])
    ids = synth.make_article_ids(2, vsize=small_cfg["V"], seed=5, max_len=24)
    with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=2 --slots=1 --flush_ms=-1")) as dec:
        for i, a in enumerate(arts):
            dec.submit(b"u%d" % i, a, b"")         # second row fills the batch: launched on slot 0, in flight
        with pytest.raises(nat.PgenError) as e:
            dec.decode_ids(*nat.pack_articles(ids, dec.max_enc_steps))
        assert e.value.code == 4
        with pytest.raises(nat.PgenError) as e:
            dec.set_option("graph=0")
        assert e.value.code == 4
        assert len(dec.poll(flush=True)) == 2
        tok, ln, lp, avg = dec.decode_ids(*nat.pack_articles(ids, dec.max_enc_steps))   # allowed again
        assert ln.tolist() == [0, 0]


@pytest.mark.parametrize("call,after", [("cudaMemcpyAsync", 0), ("cudaMemcpyAsync", 3), ("cudaMemcpyAsync", 9), ("cudaEventRecord", 0),
                                         ("cudaEventRecord", 2), ("cudaSetDevice", 0), ("cudaStreamSynchronize", 0), ("cudaLaunchKernel", 0),
                                         ("cudaLaunchKernel", 40), ("cudaGraphLaunch", 0), ("cudaStreamBeginCapture", 0)])
def test_no_row_is_lost_when_a_cuda_call_fails(fake, small_cfg, small_model_dir, call, after):
    """A CUDA call fails while a batch is being launched or collected: the call on the handle reports PGEN_ERR_CUDA, and
    once the fault is gone every row that was submitted still comes out, in order."""
    nat, L = fake
    arts = _articles(7, small_cfg["V"])
    with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=3 --slots=2 --flush_ms=-1")) as dec:
        out, errors = [], 0
        L.fake_cuda_fail(call.encode(), after)
        for i, a in enumerate(arts):
            try:
                dec.submit(b"u%d" % i, a, b"")
            except nat.PgenError as e:
                assert e.code == 3, e
                errors += 1
            try:
                out += dec.poll()
            except nat.PgenError as e:
                assert e.code == 3, e
                errors += 1
        L.fake_cuda_fail(None, 0)                  # the fault clears
        for _ in range(4):
            try:
                out += dec.poll(flush=True)
            except nat.PgenError as e:             # the injected failure may have been armed but not yet consumed
                assert e.code == 3, e
                errors += 1
        st = dec.stats()
    if call == "cudaStreamBeginCapture":           # by design not an error: the step loop falls back to plain launches
        assert errors == 0
    else:
        assert errors >= 1, "the injected failure never surfaced"
    assert [r[0] for r in out] == [b"u%d" % i for i in range(7)], (call, after, [r[0] for r in out])
    assert st["rows_in"] == st["rows_out"] == 7


def test_create_fails_cleanly_when_allocation_fails(fake, small_cfg, small_model_dir):
    nat, L = fake
    base = L.fake_cuda_live_allocs()
    for after in (0, 5, 40):
        L.fake_cuda_fail(b"cudaMalloc", after)
        with pytest.raises(nat.PgenError) as e:
            nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=4 --slots=2"))
        assert e.value.code == 3 and "cudaMalloc" in str(e.value)
        L.fake_cuda_fail(None, 0)
        assert L.fake_cuda_live_allocs() == base   # what had been allocated before the failure was freed again
    with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=4 --slots=2")) as dec:
        assert dec.flag("slots") == 2
    assert L.fake_cuda_live_allocs() == base


def test_device_ordinals(fake, small_cfg, small_model_dir):
    nat, L = fake
    assert nat.device_count() == 2
    with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=2"), device=1) as dec:
        assert dec.flag("beam_size") == 4
    with pytest.raises(nat.PgenError) as e:
        nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=2"), device=2)
    assert e.value.code == 1 and "out of range" in str(e.value)
    L.fake_cuda_set_devices(0)
    try:
        with pytest.raises(nat.PgenError) as e:
            nat.Decoder(_flags(small_model_dir, small_cfg, ""))
        assert e.value.code == 3 and "no CPU fallback" in str(e.value)
    finally:
        L.fake_cuda_set_devices(2)


def test_random_traffic_with_random_faults(fake, small_cfg, small_model_dir):
    """Random submit / poll / flush traffic, random batch and slot counts, a random CUDA call failing at a random moment
    in every round: whatever happens, rows are never lost, duplicated or reordered."""
    nat, L = fake
    rng = np.random.default_rng(2024)
    pool = _articles(40, small_cfg["V"])
    calls = [b"cudaMemcpyAsync", b"cudaEventRecord", b"cudaSetDevice", b"cudaStreamSynchronize", b"cudaLaunchKernel", b"cudaGraphLaunch"]
    for rnd in range(int(os.environ.get("PGEN_FUZZ_ROUNDS", "25"))):
        apb, slots = int(rng.integers(1, 6)), int(rng.integers(1, 4))
        with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=%d --slots=%d --flush_ms=-1" % (apb, slots))) as dec:
            n, out, sent = int(rng.integers(1, 30)), [], 0
            fault_at = int(rng.integers(0, n))
            for i in range(n):
                if i == fault_at:
                    L.fake_cuda_fail(calls[int(rng.integers(len(calls)))], int(rng.integers(0, 30)))
                try:
                    dec.submit(b"r%d-%d" % (rnd, i), pool[i % len(pool)], b"")
                except nat.PgenError as e:
                    assert e.code == 3
                sent += 1                              # a failed launch leaves the row pending: it still counts as accepted
                if rng.random() < 0.7:
                    try:
                        out += dec.poll(flush=bool(rng.random() < 0.2))
                    except nat.PgenError as e:
                        assert e.code == 3
            L.fake_cuda_fail(None, 0)
            for _ in range(3):
                try:
                    out += dec.poll(flush=True)
                except nat.PgenError as e:
                    assert e.code == 3
            assert [r[0] for r in out] == [b"r%d-%d" % (rnd, i) for i in range(sent)], (rnd, apb, slots)


def test_two_threads_two_handles(fake, small_cfg, small_model_dir):
    """Handles are independent: two threads, one handle each (the subtasks of one TaskManager), interleaved at will."""
    import threading
    nat, _ = fake
    arts = _articles(60, small_cfg["V"])
    results, errors = {}, []

    def work(k):
        try:
            with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=4 --slots=2 --flush_ms=-1"), device=k) as dec:
                out = []
                for i, a in enumerate(arts):
                    dec.submit(b"t%d-%d" % (k, i), a, b"")
                    out += dec.poll()
                out += dec.poll(flush=True)
                results[k] = [r[0] for r in out]
        except Exception as ex:                        # noqa: BLE001
            errors.append(repr(ex))
    ts = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors
    for k in range(2):
        assert results[k] == [b"t%d-%d" % (k, i) for i in range(60)]


def test_stage_mirror_deals_rows_over_subtasks_and_restores_order(fake, small_cfg, small_model_dir, monkeypatch):
    """pipeline.TFModel.transform (the Python mirror of TFModel.java:55-81) at worker_num 1, 2 and 3 on two devices: rows
    are dealt round-robin to the subtasks and come back in input order with the output schema."""
    from textsummarization_on_flink_b200 import pipeline
    nat, _ = fake
    monkeypatch.setattr(pipeline, "native", nat)
    arts = _articles(11, small_cfg["V"])
    rows = [{"uuid": "id-%d" % i, "article": a.decode(), "reference": "ref %d" % i, "extra": i} for i, a in enumerate(arts)]
    hyper = _flags(small_model_dir, small_cfg, "--articles_per_batch=2 --slots=2 --flush_ms=-1").split(" ")
    for workers in (1, 2, 3):
        m = (pipeline.TFModel().setWorkerNum(workers).setInferenceHyperParams(hyper)
             .setInferenceSelectedCols(["uuid", "article", "reference"])
             .setInferenceOutputCols(["uuid", "article", "summary", "reference"])
             .setInferenceOutputTypes([pipeline.STRING] * 4))
        out = list(m.transform(rows, devices=[0, 1]))
        assert [r["uuid"] for r in out] == [("id-%d" % i).encode() for i in range(11)]
        assert [r["article"] for r in out] == arts and all(set(r) == {"uuid", "article", "summary", "reference"} for r in out)
    m.setInferenceHyperParams(["prog", "--mode=train"])
    with pytest.raises(RuntimeError):                                  # TFModel.java:78-80: failures surface as RuntimeException
        list(m.transform(rows))


def test_tf_example_rows_and_json_records_through_the_library(fake, small_cfg, small_model_dir, monkeypatch):
    """The two wire formats either side of the stage, end to end through the C ABI on the fake runtime: serialized
    tf.Example in (pgen_submit_example) and out (pgen_example_write), and the demo's JSON Message records through
    message.transform_messages (App.java:108-131) with its max_count cut-off.  Summaries are empty here; what is checked
    is that every other field survives the trip, in order."""
    import json
    from test_tfexample_codec import Example, pb_example
    from textsummarization_on_flink_b200 import message, pipeline
    nat, _ = fake
    rows = [(b"uuid-%d" % i, b"article %d. w7 w9 . w11" % i, b"reference %d." % i) for i in range(6)]
    with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=4 --flush_ms=-1")) as dec:
        for u, a, r in rows:
            dec.submit_example(pb_example(uuid=u, article=a, reference=r))
        out = dec.poll(flush=True)
        with pytest.raises(nat.PgenError):
            dec.submit_example(pb_example(uuid=b"u", article=b"a"))        # a missing feature (batcher.py:554-557)
        assert dec.poll(flush=True) == []                                   # ... submits nothing
    assert [(u, a, r) for u, a, _, r in out] == rows
    for got in out:
        ex = Example.FromString(nat.example_write(*got))
        assert [ex.features.feature[k].bytes_list.value[0] for k in ("uuid", "article", "summary", "reference")] == list(got)
    monkeypatch.setattr(pipeline, "native", nat)
    flags = _flags(small_model_dir, small_cfg, "--articles_per_batch=4").split(" ")
    model = (pipeline.TFModel().setInferenceHyperParams(flags).setInferenceSelectedCols(["uuid", "article", "reference"])
             .setInferenceOutputCols(["uuid", "article", "summary", "reference"]).setInferenceOutputTypes([pipeline.STRING] * 4))
    recs = [message.serialize(["uuid-%d" % i, "article %d." % i, "", "reference %d." % i]) for i in range(6)]
    got = [json.loads(b) for b in message.transform_messages(model, recs, max_count=4)]
    assert [g["uuid"] for g in got] == ["uuid-%d" % i for i in range(4)]   # records 5 and 6 are past the limit
    assert all(g["article"] == "article %d." % i and g["reference"] == "reference %d." % i for i, g in enumerate(got))


def test_uploaded_copy_lists(fake, small_cfg, small_model_dir):
    """What Engine::upload puts on the device for the copy distribution (model.py:170-176: tf.scatter_nd adds the
    attention of every position onto its extended id, duplicates accumulating in position order): per article the
    distinct extended ids ascending, and for each its positions in ascending t.  Read back from the fake device."""
    nat, L = fake
    T = 24
    arts = synth.make_article_ids(5, vsize=small_cfg["V"], seed=31, max_len=T, oov_rate=0.3)
    arts.append((np.array([7, 7, 7, 0, 7], np.int32), np.array([7, 7, 7, 300, 7], np.int32), 1))
    arts.append((np.zeros(0, np.int32), np.zeros(0, np.int32), 0))
    with nat.Decoder(_flags(small_model_dir, small_cfg, "--articles_per_batch=8")) as dec:
        before = L.fake_cuda_h2d_count()
        dec.stage_upload(*nat.pack_articles(arts, T))
        assert L.fake_cuda_h2d_count() - before == 8                 # ids, ext, lens, n_oov, uniq, ucnt, pstart, plist
        n = len(arts)

        def dev(back, count):
            assert L.fake_cuda_h2d_bytes(back) == 4 * count
            return np.frombuffer(C.string_at(L.fake_cuda_h2d_dst(back), 4 * count), np.int32).copy()
        plist, pstart, ucnt, uniq = dev(0, n * T), dev(1, n * (T + 1)), dev(2, n), dev(3, n * T)
        noov, lens, ext, ids = dev(4, n), dev(5, n), dev(6, n * T), dev(7, n * T)
    for b, (a, e, k) in enumerate(arts):
        assert lens[b] == len(a) and noov[b] == k
        assert ids[b * T:b * T + len(a)].tolist() == a.tolist() and ext[b * T:b * T + len(a)].tolist() == e.tolist()
        want_u = sorted(set(e.tolist()))
        assert ucnt[b] == len(want_u) and uniq[b * T:b * T + len(want_u)].tolist() == want_u
        ps = pstart[b * (T + 1):b * (T + 1) + len(want_u) + 1].tolist()
        assert ps[0] == 0 and ps[-1] == len(a)
        for j, u in enumerate(want_u):
            pos = plist[b * T + ps[j]:b * T + ps[j + 1]].tolist()
            assert pos == [t for t in range(len(e)) if e[t] == u]     # ascending t: the order scatter_nd accumulates in


def test_launch_plan_of_a_decode_step(fake, model_dir):
    """Which kernels a full-size batch launches, in which order and with which grids -- taken from the library itself
    (the fake runtime logs every cudaLaunchKernel): 8 launches per decode step with the fused decoder cell, 13 with the
    split path, the grids DESIGN.md section 4 and profiles/r2_summary.md quote."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "fake_cuda"))
    import build_fake
    nat, L = fake
    arts = synth.make_article_ids(64, seed=3, full_len_only=True)
    plans = {}
    for mode in ("fused", "split"):
        flags = synth.hyperparams_string(model_dir, extra="--articles_per_batch=64 --slots=1 --decoder=%s" % mode)
        with nat.Decoder(flags) as dec:
            dec.set_option("graph=0")
            L.fake_cuda_log_clear()
            dec.stage_upload(*nat.pack_articles(arts, dec.max_enc_steps))
            dec.stage_run()
            dec.stage_sync()
            log = build_fake.launch_log(L)
        names = [x[0] for x in log]
        first = names.index("beam_step_kernel")
        start = max(i for i in range(first) if names[i].startswith("attn_context_kernel")) + 1    # after the step-0 attention pass
        per_step = names[start:first + 1]
        assert names[start:start + 100 * len(per_step)] == per_step * 100 and names[start + 100 * len(per_step)] == "finalize_kernel"
        plans[mode] = (per_step, {x[0]: (x[1], x[2]) for x in log})
    step, grids = plans["fused"]
    assert step == ["dec_cell_kernel", "attn_fused_kernel<4>", "attn_merge_kernel", "dec_out_kernel", "vocab_rescore_kernel<4>",
                    "vocab_tc_kernel<false>", "vocab_final_kernel", "beam_step_kernel"]
    assert grids["dec_cell_kernel"] == ((128, 1, 1), 576) and grids["attn_fused_kernel<4>"] == ((148, 1, 1), 800)
    assert grids["attn_merge_kernel"] == ((256, 1, 1), 512) and grids["dec_out_kernel"] == ((8, 16, 1), 544)
    assert grids["vocab_tc_kernel<false>"] == ((132, 1, 1), 320) and grids["vocab_final_kernel"] == ((256, 1, 1), 512)
    assert grids["beam_step_kernel"] == ((64, 1, 1), 128) and grids["enc_lstm_persistent_kernel<10>"] == ((8, 2, 7), 512)
    assert grids["gemm_tc3_kernel"] == ((148, 1, 1), 320)
    step, grids = plans["split"]
    assert len(step) == 13 and step.count("gemm_skinny_kernel<8>") + step.count("gemm_skinny_kernel<4>") == 4
    assert step[0] == "gather_step_kernel" and "attn_fused_kernel<4>" in step and "attn_pack_dfeat_kernel" not in step


def test_streaming_tool_on_the_fake_runtime(fake, model_dir):
    """tools/stream_config5.py (the BASELINE configs[4] emulation: P feeder threads of one process, one handle each) run
    as a subprocess against the fake runtime: every row comes back, rows are dealt evenly over the devices, the JSON line
    has the fields DESIGN.md section 6 quotes.  Says nothing about speed."""
    import json
    import sys
    nat, _ = fake
    env = dict(os.environ, PGEN_LIB=nat.LIB_PATH, PGEN_BENCH_MODEL=os.path.dirname(model_dir["vocab_path"]))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "stream_config5.py"), "--rows", "1500", "--parallelism", "4",
                        "--flush-ms", "20", "--pool", "64"], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["rows"] == 1500 and out["parallelism"] == 4 and out["gpus"] == 2
    assert sorted(out["rows_per_gpu"]) == [750, 750]
    assert out["latency_ms"]["p50"] <= out["latency_ms"]["p99"] <= out["latency_ms"]["max"] and 0 < out["mean_batch_fill"] <= 64
