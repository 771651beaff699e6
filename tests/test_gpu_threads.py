"""How a Flink TaskManager drives the library: several subtasks = several THREADS of one process, each with its own
handle (open -> N x flatMap -> close on its task thread; worker_num subtasks,
src/main/java/org/apache/flink/table/ml/lib/tensorflow/param/HasClusterConfig.java:20-24), and sources that deliver a
row now and then (src/test/java/.../SourceSinkTest.java:42-62: one row every 5 s).  pgen.h promises that different
handles are independent and that a lone row is not held back for ever."""
import threading
import time

import numpy as np
import pytest

import pgen_oracle as O
from conftest import small_flags
from textsummarization_on_flink_b200 import native, synth

pytestmark = pytest.mark.gpu


def test_two_threads_two_handles_one_device(small_cfg, small_model_dir):
    """Two subtask threads, each with its own handle on the same GPU, decoding different rows at the same time: both
    get exactly the oracle's rows, in their own submission order (ctypes releases the GIL inside every call)."""
    hps_flags = small_flags(small_model_dir, small_cfg, extra="--articles_per_batch=4 --flush_ms=-1 --slots=2")
    hps = O.parse_hyperparams(hps_flags)
    vocab = O.Vocab(small_model_dir["vocab_path"], small_cfg["V"])
    results, errors = {}, []
    start = threading.Barrier(2)

    def subtask(k):
        try:
            rows = synth.make_articles(22, vsize=small_cfg["V"], seed=300 + k, max_len=24, oov_rate=0.2)
            with native.Decoder(hps_flags) as dec:
                start.wait()
                got = []
                for rep in range(3):                                   # 66 rows per subtask
                    for i, a in enumerate(rows):
                        dec.submit(b"t%d-%d-%d" % (k, rep, i), a, b"")
                        got += dec.poll()
                got += dec.poll(flush=True)
            results[k] = (rows, got)
        except Exception as ex:                                        # surfaces in the main thread
            errors.append(ex)

    threads = [threading.Thread(target=subtask, args=(k,)) for k in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(120)
    assert not errors, errors
    for k in range(2):
        rows, got = results[k]
        assert [g[0] for g in got] == [b"t%d-%d-%d" % (k, rep, i) for rep in range(3) for i in range(len(rows))]
        ref = [O.decode_row(small_cfg["W"], vocab, hps, b"", a, b"", mode="dedup")[2] for a in rows]
        for j, g in enumerate(got):
            assert g[2] == ref[j % len(rows)], "subtask %d row %d" % (k, j)


def test_lone_row_comes_back_after_the_flush_deadline(small_cfg, small_model_dir):
    """One row, then silence (SourceSinkTest's cadence): with --flush_ms=50 the row is decoded by a later poll without
    any further submit and without flush=1 -- the reference held such rows until the next one arrived
    (doc/Flink-AI-Extended Integration Report.md:341-376)."""
    flags = small_flags(small_model_dir, small_cfg, extra="--articles_per_batch=64 --flush_ms=50")
    hps = O.parse_hyperparams(flags)
    vocab = O.Vocab(small_model_dir["vocab_path"], small_cfg["V"])
    article = synth.make_articles(1, vsize=small_cfg["V"], seed=9, max_len=24)[0]
    with native.Decoder(flags) as dec:
        dec.submit(b"lone", article, b"r")
        assert dec.poll() == []                                        # deadline not reached: the row waits for company
        t0 = time.time()
        got = []
        while not got and time.time() - t0 < 5.0:
            time.sleep(0.02)
            got = dec.poll()
        waited = time.time() - t0
    assert len(got) == 1 and got[0] == O.decode_row(small_cfg["W"], vocab, hps, b"lone", article, b"r", mode="dedup")
    assert waited < 2.0, "the lone row took %.2f s with a 50 ms flush deadline" % waited


def test_stage_api_is_refused_while_rows_are_in_flight(small_cfg, small_model_dir):
    """pgen_stage_* / pgen_decode_ids address a slot's engine directly; with a pgen_submit batch in flight on that slot
    they must fail (PGEN_ERR_STATE) instead of overwriting its buffers, and work again after the rows were collected."""
    flags = small_flags(small_model_dir, small_cfg, extra="--articles_per_batch=4 --flush_ms=-1 --slots=1")
    rows = synth.make_articles(4, vsize=small_cfg["V"], seed=12, max_len=24)
    ids = synth.make_article_ids(3, vsize=small_cfg["V"], seed=13, max_len=24)
    with native.Decoder(flags) as dec:
        for i, a in enumerate(rows):
            dec.submit(b"r%d" % i, a, b"")                             # the 4th submit launches the batch on slot 0
        with pytest.raises(native.PgenError) as e:
            dec.decode_ids(*native.pack_articles(ids, dec.max_enc_steps))
        assert e.value.code == 4 and "in flight" in str(e.value)
        with pytest.raises(native.PgenError) as e:
            dec.set_option("graph=0")
        assert e.value.code == 4
        got = dec.poll(flush=True)
        assert [g[0] for g in got] == [b"r%d" % i for i in range(4)]
        tok, ln, lp, avg = dec.decode_ids(*native.pack_articles(ids, dec.max_enc_steps))
        assert (ln > 0).all()
        assert dec.stats()["rows_in"] == dec.stats()["rows_out"] == 4


def test_device_count_export():
    import torch
    assert native.device_count() == torch.cuda.device_count() >= 1


def test_create_refuses_sizes_beyond_shared_memory(small_cfg, small_model_dir):
    """max_enc_steps so large that the softmax / selection kernels cannot hold a row in shared memory: pgen_create
    says so (PGEN_ERR_UNSUPPORTED with the numbers) instead of failing at the first batch with a launch error."""
    flags = small_flags(small_model_dir, small_cfg, beam=8, max_enc=20000, extra="--articles_per_batch=1 --slots=1")
    with pytest.raises(native.PgenError) as e:
        native.Decoder(flags)
    assert e.value.code == 5 and "shared memory" in str(e.value)
