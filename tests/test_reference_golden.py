"""Golden vectors produced by the reference's OWN Python sources (tests/golden/ref_small.npz, ref_full.npz; generated
here, where /root/reference exists, by tests/golden/make_reference_golden.py: model.py, attention_decoder.py,
beam_search.py, decode.py, batcher.py and data.py of the reference executing on oracle/tf1_shim.py, a NumPy restatement
of the TensorFlow 1.11 ops they call).

CPU: (1) those vectors equal the oracle-made vectors the GPU tests compare with (decode_small.npz, decode_full.npz) --
identical token ids, log-probs within 1e-5 (300-word model) / 5e-4 (50 000-word model, 400 encoder + 100 decoder
steps); (2) the oracle reproduces the reference run step by step (every top-2K id and log-prob of every beam row of
every decode step) and row by row (ids, OOV lists, summary strings); (3) the TF variable names and shapes the
library's weight archive uses are the ones the reference's graph creates; (4) the shim's ops agree with PyTorch's
independent implementations of the same ops.  Where /root/reference exists the generator is re-run and must reproduce
the committed file."""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch

import pgen_oracle as O
import tf1_shim as S
from textsummarization_on_flink_b200 import synth

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(HERE, name + ".py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


mg = _load("make_golden")


def _small_arts():
    c = mg.SMALL
    return synth.make_article_ids(c["n"], vsize=c["V"], seed=c["art_seed"], max_len=c["max_len"], oov_rate=c["oov_rate"])


def test_reference_run_equals_the_small_golden_vectors():
    g, r = np.load(os.path.join(HERE, "decode_small.npz")), np.load(os.path.join(HERE, "ref_small.npz"))
    assert np.array_equal(r["lens"], g["lens"]) and np.array_equal(r["tokens"], g["tokens"])
    np.testing.assert_allclose(r["logprobs"], g["logprobs"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(r["avg"], g["avg"], rtol=0, atol=1e-5)
    assert (g["lens"] < 30).sum() >= 6 and (g["lens"] == 30).sum() >= 2      # early [STOP]s and full-length summaries


def test_reference_run_equals_the_full_size_golden_vectors():
    g, r = np.load(os.path.join(HERE, "decode_full.npz")), np.load(os.path.join(HERE, "ref_full.npz"))
    assert np.array_equal(r["lens"], g["lens"]) and np.array_equal(r["tokens"], g["tokens"])
    np.testing.assert_allclose(r["logprobs"], g["logprobs"], rtol=5e-4, atol=1e-5)
    np.testing.assert_allclose(r["avg"], g["avg"], rtol=5e-4, atol=1e-5)
    assert r["tokens"].shape == (4, 100) and int(r["lens"].min()) >= 35


def test_oracle_equals_the_reference_run_on_reference_scale_weights():
    """BASELINE.json configs[0] ("random-init weights"): W0 = the reference's own initialisers at full size.  The
    vocabulary distribution is then almost flat and the copy distribution decides; one article in the fast oracle mode."""
    r = np.load(os.path.join(HERE, "ref_full.npz"))
    a, e, k = synth.make_article_ids(2, seed=4242)[1]
    hps = O.parse_hyperparams(mg.FULL["flags"])
    t, l, av = O.decode_ids(synth.make_weights("W0"), hps, a, e, k, mode="dedup")
    n = int(r["w0_lens"][1])
    assert t == r["w0_tokens"][1, :n].tolist()
    np.testing.assert_allclose(np.array(l, np.float32), r["w0_logprobs"][1, :n], rtol=1e-4, atol=1e-5)


def test_oracle_follows_the_reference_run_step_by_step(monkeypatch):
    """Every decode_onestep of the reference run (model.py:367-443: top-2K ids and log-probs of all beam rows) against
    the oracle's decode_onestep on the same article, for all 12 articles of the small case."""
    r = np.load(os.path.join(HERE, "ref_small.npz"))
    W = mg.small_weights()
    hps = O.parse_hyperparams(mg.SMALL["flags"])
    steps = []
    real = O.decode_onestep

    def spy(*a, **k):
        out = real(*a, **k)
        steps.append((np.array(out[0]), np.array(out[1])))
        return out
    monkeypatch.setattr(O, "decode_onestep", spy)
    n_steps = 0
    for b, (a, e, k) in enumerate(_small_arts()):
        steps.clear()
        toks, lps, avg = O.decode_ids(W, hps, a, e, k, mode="literal")
        assert len(steps) == int(r["n_steps"][b])
        for s, (ids, lp) in enumerate(steps):
            want_ids, want_lp = r["step_ids"][b, s], r["step_logp"][b, s]
            fin = np.isfinite(want_lp)
            assert np.array_equal(np.isfinite(lp), fin)
            np.testing.assert_allclose(lp[fin], want_lp[fin], rtol=0, atol=2e-5)
            diff = ids != want_ids
            if diff.any():                              # only where two candidates of a row tie within the tolerance
                for i, j in zip(*np.nonzero(diff)):
                    assert abs(float(want_lp[i, j]) - float(lp[i, j])) <= 2e-5 or not fin[i, j]
            n_steps += 1
        assert toks == r["tokens"][b, :r["lens"][b]].tolist()
        assert abs(avg - float(r["avg"][b])) < 1e-5
    assert n_steps > 150


def test_oracle_text_path_equals_the_reference_run(tmp_path):
    """data.py / batcher.py / decode.py of the reference on article TEXT: encoder ids, extended ids, OOV lists, and the
    (uuid, article, summary, reference) rows its writer received."""
    r = np.load(os.path.join(HERE, "ref_small.npz"))
    c = mg.SMALL
    vpath = str(tmp_path / "vocab")
    synth.write_vocab_file(vpath, c["V"])
    vocab = O.Vocab(vpath, c["V"])
    W = mg.small_weights()
    hps = O.parse_hyperparams(mg.SMALL["flags"])
    arts = _small_arts()
    some_oov = 0
    for i, (ids, ext, n_oov) in enumerate(arts):
        text = synth.ids_to_article(ext, c["V"], salt=i)
        assert str(r["row_article"][i]).encode() == text and str(r["row_uuid"][i]) == "uuid-%d" % i
        ex = O.make_example(text, vocab, 24)
        n = int(r["enc_lens"][i])
        assert len(ex.enc_input) == n == len(ids)
        assert ex.enc_input == r["enc_batch"][i, :n].tolist() == ids.tolist()
        assert ex.enc_input_extend_vocab == r["enc_batch_extend_vocab"][i, :n].tolist() == ext.tolist()
        oovs = [w.encode() for w in str(r["art_oovs"][i]).split("\t") if w]
        assert ex.article_oovs == oovs and len(oovs) == int(r["n_oovs"][i]) == n_oov
        some_oov += n_oov > 0
        row = O.decode_row(W, vocab, hps, b"uuid-%d" % i, text, str(r["in_reference"][i]).encode())
        assert row[0].decode() == str(r["row_uuid"][i]) and row[1].decode() == str(r["row_article"][i])
        assert row[2].decode() == str(r["row_summary"][i])
        # the reference column: the reference re-tokenises it (nltk) and joins the sentences with one space; on an
        # already tokenised reference that is the identity, which is what the library's pass-through assumes
        assert row[3].decode() == str(r["row_reference"][i]) == str(r["in_reference"][i])
    assert some_oov >= 6
    assert any("  " in str(s) for s in r["row_summary"])          # sentences joined by two spaces (decode.py:182)


@pytest.mark.parametrize("beam,cov,max_dec,min_dec", [(8, 1, 20, 5), (3, 0, 16, 4), (1, 1, 12, 3), (2, 0, 12, 12)])
@pytest.mark.parametrize("mode", ["literal", "dedup"])
def test_oracle_equals_the_reference_run_across_beam_sizes_and_coverage(beam, cov, max_dec, min_dec, mode):
    """ref_variants.npz: beam 8 / 3 / 1 / 2, with and without coverage (a graph without coverage has no w_c variable),
    min_dec_steps == max_dec_steps (no [STOP] is ever accepted).  Both oracle modes: `literal` does the reference's
    work, `dedup` is the arithmetic the CUDA path implements (features once, one attention pass per step)."""
    v = np.load(os.path.join(HERE, "ref_variants.npz"))
    tag = "b%dc%d_" % (beam, cov)
    c = mg.SMALL
    W = mg.small_weights()
    hps = O.parse_hyperparams("p --mode=decode --beam_size=%d --coverage=%d --max_enc_steps=24 --max_dec_steps=%d --min_dec_steps=%d"
                              % (beam, cov, max_dec, min_dec))
    arts = synth.make_article_ids(5, vsize=c["V"], seed=321, max_len=c["max_len"], oov_rate=0.25)
    for b, (a, e, k) in enumerate(arts):
        toks, lps, avg = O.decode_ids(W, hps, a, e, k, mode=mode)
        n = int(v[tag + "lens"][b])
        assert toks == v[tag + "tokens"][b, :n].tolist()
        np.testing.assert_allclose(np.array(lps, np.float32), v[tag + "logprobs"][b, :n], rtol=0, atol=2e-5)
        assert abs(avg - float(v[tag + "avg"][b])) < 1e-5
        assert int(v[tag + "n_steps"][b]) <= max_dec


@pytest.mark.parametrize("mode", ["literal", "dedup"])
def test_oracle_equals_the_reference_run_without_pointer_gen_and_on_exact_ties(mode):
    """The baseline model (pointer_gen off: no p_gen layer, no copy distribution, OOVs stay [UNK]) and an all-zero model
    on which every vocabulary word ties exactly, so that only the rules decide: tf.nn.top_k prefers the lower id,
    sort_hyps keeps equal hypotheses in their original order (beam_search.py:171-173)."""
    v = np.load(os.path.join(HERE, "ref_variants.npz"))
    c = mg.SMALL
    arts = synth.make_article_ids(5, vsize=c["V"], seed=321, max_len=c["max_len"], oov_rate=0.25)
    zero = {k: np.zeros_like(x) for k, x in mg.small_weights().items()}
    cases = [("nopg_", mg.small_weights(), "p --mode=decode --beam_size=4 --coverage=1 --pointer_gen=0 --max_enc_steps=24 --max_dec_steps=14 --min_dec_steps=4"),
             ("zero_", zero, "p --mode=decode --beam_size=4 --coverage=1 --max_enc_steps=24 --max_dec_steps=8 --min_dec_steps=3")]
    for tag, W, flags in cases:
        hps = O.parse_hyperparams(flags)
        for b, (a, e, k) in enumerate(arts):
            if tag == "nopg_":
                e, k = a, 0                                            # batcher.py:66-76: no extended ids without pointer_gen
            toks, lps, avg = O.decode_ids(W, hps, a, e, k, mode=mode)
            n = int(v[tag + "lens"][b])
            assert toks == v[tag + "tokens"][b, :n].tolist(), (tag, b)
            np.testing.assert_allclose(np.array(lps, np.float32), v[tag + "logprobs"][b, :n], rtol=0, atol=2e-5)
            assert abs(avg - float(v[tag + "avg"][b])) < 1e-5
    assert (v["zero_step_logp"][0, 0, 0, 1:4] == v["zero_step_logp"][0, 0, 0, 1]).all()      # exact ties in the fixture
    assert (np.diff(v["zero_step_ids"][0, 0, 0, 1:4]) > 0).all()                              # ... resolved by id


def test_oracle_text_edges_equal_the_reference_run(tmp_path):
    """Truncation at max_enc_steps, special tokens spelled out in an article, repeated OOVs, one-word and all-OOV articles,
    blanks / tabs / newlines / vertical tabs / form feeds between words: ids, OOV lists, decoded summaries."""
    v = np.load(os.path.join(HERE, "ref_variants.npz"))
    c = mg.SMALL
    vpath = str(tmp_path / "vocab")
    synth.write_vocab_file(vpath, c["V"])
    vocab = O.Vocab(vpath, c["V"])
    W = mg.small_weights()
    hps = O.parse_hyperparams(mg.SMALL["flags"])
    assert int(v["edge_enc_lens"][0]) == 24 and len(str(v["edge_article"][0]).split()) == 40          # truncated (batcher.py:55-57)
    assert str(v["edge_oovs"][1]) == "[UNK]\tzz" and str(v["edge_enc"][1]).split()[:2] == ["3", "0"]    # "[STOP]" is id 3, "[UNK]" an OOV
    for i, art in enumerate(v["edge_article"]):
        text = str(art).encode()
        ex = O.make_example(text, vocab, 24)
        assert ex.enc_input == [int(x) for x in str(v["edge_enc"][i]).split()]
        assert ex.enc_input_extend_vocab == [int(x) for x in str(v["edge_ext"][i]).split()]
        assert ex.article_oovs == [w.encode() for w in str(v["edge_oovs"][i]).split("\t") if w]
        row = O.decode_row(W, vocab, hps, b"e-%d" % i, text, b"edge reference %d ." % i)
        assert row[2].decode() == str(v["edge_summary"][i])
        toks, lps, _ = O.decode_ids(W, hps, ex.enc_input, ex.enc_input_extend_vocab, len(ex.article_oovs), mode="dedup")
        n = int(v["edge_lens"][i])
        assert toks == v["edge_tokens"][i, :n].tolist()
        np.testing.assert_allclose(np.array(lps, np.float32), v["edge_logprobs"][i, :n], rtol=0, atol=2e-5)


def test_library_text_path_equals_the_reference_run(tmp_path):
    """The PRODUCT's host text path (csrc/text.cc through the C ABI, pgen_text_open / _encode_article / _decode_summary:
    no GPU needed) against what the reference's data.py / batcher.py / decode.py produced: encoder ids, extended ids, OOV
    counts, and the summary string of the tokens the reference's beam search returned."""
    from textsummarization_on_flink_b200 import native
    r, v = np.load(os.path.join(HERE, "ref_small.npz")), np.load(os.path.join(HERE, "ref_variants.npz"))
    c = mg.SMALL
    vpath = str(tmp_path / "vocab")
    synth.write_vocab_file(vpath, c["V"])
    with native.TextPath(vpath, c["V"], 24) as tp:
        assert tp.vocab_size() == c["V"]
        for i in range(len(r["row_article"])):
            ids, ext, n_oov = tp.encode(str(r["row_article"][i]).encode())
            n = int(r["enc_lens"][i])
            assert ids.tolist() == r["enc_batch"][i, :n].tolist() and ext.tolist() == r["enc_batch_extend_vocab"][i, :n].tolist()
            assert n_oov == int(r["n_oovs"][i])
            assert tp.summary(r["tokens"][i, :r["lens"][i]]).decode() == str(r["row_summary"][i])
        for i, art in enumerate(v["edge_article"]):
            ids, ext, n_oov = tp.encode(str(art).encode())
            assert ids.tolist() == [int(x) for x in str(v["edge_enc"][i]).split()]
            assert ext.tolist() == [int(x) for x in str(v["edge_ext"][i]).split()]
            assert n_oov == len([w for w in str(v["edge_oovs"][i]).split("\t") if w])
            assert tp.summary(v["edge_tokens"][i, :v["edge_lens"][i]]).decode() == str(v["edge_summary"][i])
        # summaries of the other reference runs (same five articles): beam 8 / 3 / 1 / 2 and the all-zero model
        arts = synth.make_article_ids(5, vsize=c["V"], seed=321, max_len=c["max_len"], oov_rate=0.25)
        for tag in ("b8c1_", "b3c0_", "b1c1_", "b2c0_", "zero_"):
            for b, (_, ext, _) in enumerate(arts):
                tp.encode(synth.ids_to_article(ext, c["V"], salt=b))
                assert tp.summary(v[tag + "tokens"][b, :v[tag + "lens"][b]]).decode() == str(v[tag + "summary"][b]), (tag, b)


def test_library_text_path_equals_the_oracle_on_odd_input(tmp_path):
    """csrc/text.cc against the (reference-pinned) oracle where the committed vectors do not reach: random articles with
    many OOVs, empty and blank articles, every ASCII white-space byte, bytes >= 0x80 (Python 2 str.split leaves them
    alone), random token lists with and without [STOP], ids outside the extended vocabulary."""
    from textsummarization_on_flink_b200 import native
    V = 300
    vpath = str(tmp_path / "vocab")
    synth.write_vocab_file(vpath, V)
    vocab = O.Vocab(vpath, V)
    rng = np.random.default_rng(0)
    arts = synth.make_articles(40, vsize=V, seed=4, max_len=40, oov_rate=0.3)
    arts += [b"", b" \t\n", b"article 0.", b"a\tb\nc\rd\x0be\x0cf  g", b"[STOP] [UNK] . , w7 w7 zz zz", b"caf\xc3\xa9 \xa0x\x85y \xff w9",
             b"w8\x1cw9 \x1d \x1e\x1f"]
    with native.TextPath(vpath, V, 24) as tp:
        for a in arts:
            ids, ext, n_oov = tp.encode(a)
            ex = O.make_example(a, vocab, 24)
            assert ids.tolist() == ex.enc_input and ext.tolist() == ex.enc_input_extend_vocab, a
            assert n_oov == len(ex.article_oovs)
            hi = V + n_oov
            toks = [t if t != 3 else 4 for t in rng.integers(0, hi, size=30).tolist()]
            for cut in (toks, toks[:7] + [3] + toks[7:], [3], [], [4, 4, 3], [4, 7, 4, 4, 9]):
                want = O.format_summary(O.cut_at_stop(O.outputids2words(cut, vocab, ex.article_oovs)))
                assert tp.summary(cut) == want
            with pytest.raises(native.PgenError):
                tp.summary([hi])
    with pytest.raises(native.PgenError) as e:
        native.TextPath(str(tmp_path / "missing"), V, 24)
    assert e.value.code == 2
    # many distinct OOVs in one article (BASELINE configs[3]: 200), repeated in random order; long words that share their
    # first eight bytes with vocabulary words and with each other
    many = synth.make_articles(6, vsize=V, seed=8, max_len=400, fixed_oovs=200)
    many.append(b" ".join(b"prefix__%d" % (i % 37) for i in range(300)) + b" w12345678 w1234567 w123456789 w12")
    with native.TextPath(vpath, V, 400) as tp:
        for a in many:
            ids, ext, n_oov = tp.encode(a)
            ex = O.make_example(a, vocab, 400)
            assert ids.tolist() == ex.enc_input and ext.tolist() == ex.enc_input_extend_vocab and n_oov == len(ex.article_oovs)
            toks = list(range(V, V + n_oov))
            assert tp.summary(toks) == O.format_summary(O.outputids2words(toks, vocab, ex.article_oovs))
    assert n_oov == 37 + 3
    long_vocab = str(tmp_path / "long_vocab")
    with open(long_vocab, "w") as f:
        f.write("".join("%s 1\n" % w for w in ("abcdefgh", "abcdefghi", "abcdefghj", "abcdefghij", "abcdefg", "x" * 40, "x" * 39 + "y")))
    lv = O.Vocab(long_vocab, 0)
    with native.TextPath(long_vocab, 0, 50) as tp:
        a = b"abcdefgh abcdefghi abcdefghj abcdefghij abcdefg abcdefghk abcdefghijk " + b"x" * 40 + b" " + b"x" * 39 + b"y " + b"x" * 39 + b"z"
        ids, ext, n_oov = tp.encode(a)
        ex = O.make_example(a, lv, 50)
        assert ids.tolist() == ex.enc_input == [4, 5, 6, 7, 8, 0, 0, 9, 10, 0] and ext.tolist() == ex.enc_input_extend_vocab and n_oov == 3


def test_vocabulary_files_load_like_the_reference_loads_them(tmp_path):
    """data.py:40-75 run on twelve crafted vocabulary files (max_size reached, fewer than the specials, malformed /
    empty lines, tabs, duplicates, reserved tokens): the library's loader (csrc/text.cc through pgen_text_open) and the
    oracle's agree with the reference on the resulting vocabulary, and fail where it raises -- the library with the
    reference's own message."""
    from textsummarization_on_flink_b200 import native
    v = np.load(os.path.join(HERE, "ref_variants.npz"))
    assert len(v["vocab_file"]) == 12
    for i, (text, max_size, outcome) in enumerate(zip(v["vocab_file"], v["vocab_max"], v["vocab_outcome"])):
        path = str(tmp_path / ("vocab%d" % i))
        with open(path, "w", newline="") as f:
            f.write(str(text))
        kind, rest = str(outcome).split("\t", 1)
        if kind == "error":
            with pytest.raises(native.PgenError) as e:
                native.TextPath(path, int(max_size), 10)
            assert e.value.code == 2 and str(e.value).endswith(rest), (str(e.value), rest)
            with pytest.raises(ValueError):
                O.Vocab(path, int(max_size))
            continue
        size, words = rest.split("\t")
        words = words.split(" ")
        ov = O.Vocab(path, int(max_size))
        assert ov.size() == int(size) and [ov.id2word(j).decode() for j in range(ov.size())] == words
        with native.TextPath(path, int(max_size), 10) as tp:
            assert tp.vocab_size() == int(size)
            tp.encode(b"")
            for j, w in enumerate(words):
                if j != 3:                                             # [STOP] ends a summary: nothing is printed for it
                    assert tp.summary([j]).decode() == w
            ids, ext, n_oov = tp.encode(" ".join(words[4:] + ["zzz"]).encode())
            assert ids.tolist() == list(range(4, int(size))) + [0] and ext.tolist()[-1] == int(size) and n_oov == 1


def test_beam_bookkeeping_equals_the_references_on_exact_ties(monkeypatch):
    """160 cases of the reference's run_beam_search / Hypothesis / sort_hyps (beam_search.py, run by the generator) on a
    deterministic stand-in for decode_onestep whose log-probs are multiples of 0.25: hypotheses tie exactly, so the
    stable descending sort, the [STOP] / min_dec_steps rule, the single parent of step 0, [UNK] for OOV ids and the
    results-or-live-hypotheses ending are all that decides.  The oracle's beam search must return the same hypothesis."""
    bf = _load("beam_fuzz")
    v = np.load(os.path.join(HERE, "ref_variants.npz"))
    assert len(bf.CASES) == len(v["beam_tokens"]) == 160
    ends = {"stop": 0, "live": 0}
    for n, case in enumerate(bf.CASES):
        K, V, T = case["K"], case["V"], 5
        W = {O.N_EMB: np.zeros((V, 2), np.float32)}
        hps = O.parse_hyperparams("p --mode=decode --beam_size=%d --coverage=1 --max_dec_steps=%d --min_dec_steps=%d"
                                  % (K, case["max_dec"], case["min_dec"]))
        step = [0]

        def fake_encoder(W_, ids):
            return np.zeros((T, 2), np.float32), (np.zeros(1, np.float32), np.zeros(1, np.float32))

        def fake_onestep(W_, hps_, enc_states, feat, mask, ext, n_oov, latest, c, h, cov):
            assert all(0 <= t < V for t in latest)
            ids, logp = bf.fake_step(case, step[0])
            step[0] += 1
            R = len(latest)
            return ids, logp, np.zeros((R, 1), np.float32), np.zeros((R, 1), np.float32), np.full((R, T), 0.2, np.float32), \
                np.full((R,), 0.5, np.float32), np.zeros((R, T), np.float32), None
        monkeypatch.setattr(O, "run_encoder", fake_encoder)
        monkeypatch.setattr(O, "decode_onestep", fake_onestep)
        best = O.run_beam_search(W, hps, [4] * T, [4] * T, case["n_oov"], mode="literal")
        want = [int(t) for t in v["beam_tokens"][n] if t >= 0]
        assert [int(t) for t in best.tokens] == want, (n, case)
        np.testing.assert_allclose([float(x) for x in best.log_probs], v["beam_logp"][n][:len(want)], rtol=0, atol=0)
        assert best.avg_log_prob == float(v["beam_avg"][n])
        ends["stop" if want[-1] == 3 else "live"] += 1
    assert ends["stop"] >= 40 and ends["live"] >= 40


def test_flag_table_is_the_references():
    """The 27 tf.app.flags.DEFINE_* statements of run_summarization.py:47-87, executed by the generator: names, kinds and
    defaults against the oracle's table and against the library's parser (pgen_parse_flag: no device needed)."""
    import ast
    from textsummarization_on_flink_b200 import native
    v = np.load(os.path.join(HERE, "ref_variants.npz"))
    ref = {str(n): (str(k), ast.literal_eval(str(d))) for k, n, d in zip(v["flag_kind"], v["flag_name"], v["flag_default"])}
    assert len(ref) == 27
    kinds = {"string": str, "integer": int, "boolean": bool, "float": float}
    assert {n: (kinds[k], d) for n, (k, d) in ref.items()} == O._FLAG_DEFAULTS
    for name, (kind, default) in ref.items():
        if kind in ("integer", "boolean"):
            assert native.parse_flag("prog", name) == int(default), name
            if kind == "boolean":
                assert native.parse_flag("prog --%s" % name, name) == 1 and native.parse_flag("prog --no%s" % name, name) == 0
                assert native.parse_flag("prog --%s=false" % name, name) == 0 and native.parse_flag("prog --%s=1" % name, name) == 1
            else:
                assert native.parse_flag("prog --%s=7 --unknown_flag=3" % name, name) == 7
        else:
            with pytest.raises(native.PgenError):
                native.parse_flag("prog", name)
    # default_setup (run_summarization.py:312-313): decode mode overwrites batch_size with beam_size; token 0 is a program name
    assert native.parse_flag("prog --mode=decode --batch_size=16 --beam_size=5", "batch_size") == 5
    assert native.parse_flag("prog --mode=train --batch_size=16 --beam_size=5", "batch_size") == 16
    assert native.parse_flag("--beam_size=9 --beam_size=6", "beam_size") == 6
    assert O.parse_hyperparams("prog --mode=decode --batch_size=16 --beam_size=5")["batch_size"] == 5
    with pytest.raises(native.PgenError):
        native.parse_flag("prog --beam_size=x", "beam_size")


@pytest.mark.parametrize("name,V,H,E", [("ref_small.npz", 300, 32, 16), ("ref_full.npz", 50000, 256, 128)])
def test_weight_archive_names_are_the_reference_graphs_variables(name, V, H, E):
    """The reference's graph (built by its own model.py / attention_decoder.py) creates exactly the variables the .pgenw
    archive is keyed by, with the same shapes: the generator's load_ckpt stand-in fails on any difference, and the
    result is recorded in the vectors."""
    r = np.load(os.path.join(HERE, name))
    got = {str(n): tuple(int(x) for x in str(s).split(",")) for n, s in zip(r["variable_names"], r["variable_shapes"])}
    assert got == {k: tuple(v) for k, v in synth.variable_shapes(V, H, E).items()}
    assert got == {k: tuple(v) for k, v in O.variable_shapes(V, H, E).items()}


# ---- the shim's ops against PyTorch's implementations of the same ops ----
def _run(t, feed=None):
    return S.Session().run(t, feed or {})


def test_shim_lstm_and_bidirectional_rnn_match_torch():
    S.reset_default_graph()
    rng = np.random.default_rng(0)
    B, T, E, H = 3, 7, 5, 4
    x = rng.standard_normal((B, T, E)).astype(np.float32)
    lens = np.array([7, 4, 1], np.int32)
    cf, cb = S.LSTMCell(H), S.LSTMCell(H)
    with S.variable_scope("encoder"):
        (of, ob), (sf, sb) = S.bidirectional_dynamic_rnn(cf, cb, S.convert_to_tensor(x), sequence_length=S.convert_to_tensor(lens))
    names = [v.var_name for v in S.global_variables()]
    assert names == ["encoder/bidirectional_rnn/fw/lstm_cell/kernel", "encoder/bidirectional_rnn/fw/lstm_cell/bias",
                     "encoder/bidirectional_rnn/bw/lstm_cell/kernel", "encoder/bidirectional_rnn/bw/lstm_cell/bias"]
    for v in S.global_variables():
        v.assign_value(rng.uniform(-0.5, 0.5, size=v.value.shape).astype(np.float32))
    got_f, got_b, (cfin, hfin), (cbin, hbin) = _run([of, ob, sf, sb])
    lstm = torch.nn.LSTM(E, H, batch_first=True, bidirectional=True)
    perm = np.concatenate([np.arange(0, H), np.arange(2 * H, 3 * H), np.arange(H, 2 * H), np.arange(3 * H, 4 * H)])   # TF i,j,f,o -> torch i,f,g,o
    with torch.no_grad():
        for cell, sfx in ((cf, ""), (cb, "_reverse")):
            k, b = cell._kernel.value, cell._bias.value.copy()
            b[2 * H:3 * H] += 1.0                                      # forget_bias
            getattr(lstm, "weight_ih_l0" + sfx).copy_(torch.from_numpy(k[:E].T[perm].copy()))
            getattr(lstm, "weight_hh_l0" + sfx).copy_(torch.from_numpy(k[E:].T[perm].copy()))
            getattr(lstm, "bias_ih_l0" + sfx).copy_(torch.from_numpy(b[perm].copy()))
            getattr(lstm, "bias_hh_l0" + sfx).zero_()
        packed = torch.nn.utils.rnn.pack_padded_sequence(torch.from_numpy(x), torch.from_numpy(lens.astype(np.int64)), batch_first=True)
        out, (hn, cn) = lstm(packed)
        out, _ = torch.nn.utils.rnn.pad_packed_sequence(out, batch_first=True, total_length=T)
    np.testing.assert_allclose(np.concatenate([got_f, got_b], axis=2), out.numpy(), atol=2e-6)
    np.testing.assert_allclose(hfin, hn[0].numpy(), atol=2e-6)
    np.testing.assert_allclose(cfin, cn[0].numpy(), atol=2e-6)
    np.testing.assert_allclose(hbin, hn[1].numpy(), atol=2e-6)
    np.testing.assert_allclose(cbin, cn[1].numpy(), atol=2e-6)


def test_shim_pointwise_ops_match_torch():
    S.reset_default_graph()
    rng = np.random.default_rng(1)
    x = rng.standard_normal((4, 9)).astype(np.float32) * 3
    np.testing.assert_allclose(_run(S.softmax(S.convert_to_tensor(x))), torch.softmax(torch.from_numpy(x), -1).numpy(), atol=1e-7)
    np.testing.assert_allclose(_run(S.sigmoid(S.convert_to_tensor(x))), torch.sigmoid(torch.from_numpy(x)).numpy(), atol=1e-7)
    # top_k: descending, lower index first among equal values
    y = np.array([[1., 3., 3., 0., 3.], [2., 2., 2., 2., 2.]], np.float32)
    v, i = _run(list(S.top_k(S.convert_to_tensor(y), 4)))
    assert i.tolist() == [[1, 2, 4, 0], [0, 1, 2, 3]] and v.tolist() == [[3, 3, 3, 1], [2, 2, 2, 2]]
    # scatter_nd: duplicates accumulate
    idx = np.array([[[0, 2], [0, 2], [0, 0]], [[1, 1], [1, 3], [1, 1]]], np.int32)
    upd = np.array([[1., 2., 4.], [8., 16., 32.]], np.float32)
    got = _run(S.scatter_nd(S.convert_to_tensor(idx), S.convert_to_tensor(upd), [2, S.convert_to_tensor(np.int32(4))]))
    want = torch.zeros(2, 4).index_put_((torch.from_numpy(idx[..., 0].astype(np.int64)), torch.from_numpy(idx[..., 1].astype(np.int64))),
                                        torch.from_numpy(upd), accumulate=True)
    assert got.tolist() == want.tolist() == [[4, 0, 3, 0], [0, 40, 0, 16]]
    # conv2d with a 1 x 1 filter, NHWC
    a = rng.standard_normal((2, 6, 1, 5)).astype(np.float32)
    w = rng.standard_normal((1, 1, 5, 3)).astype(np.float32)
    got = _run(S.conv2d(S.convert_to_tensor(a), S.convert_to_tensor(w), [1, 1, 1, 1], "SAME"))
    want = torch.nn.functional.conv2d(torch.from_numpy(a).permute(0, 3, 1, 2), torch.from_numpy(w).permute(3, 2, 0, 1)).permute(0, 2, 3, 1)
    np.testing.assert_allclose(got, want.numpy(), atol=1e-5)


def test_shim_variable_scope_rules():
    S.reset_default_graph()
    with S.variable_scope("a") as sc:
        v = S.get_variable("w", [2])
        with pytest.raises(ValueError):
            S.get_variable("w", [2])                                   # duplicate without reuse
        with S.variable_scope(S.get_variable_scope(), reuse=True):
            assert S.get_variable("w", [2]) is v
            with S.variable_scope("inner"):                            # reuse is inherited
                with pytest.raises(ValueError):
                    S.get_variable("missing", [1])
        S.get_variable_scope().reuse_variables()
        assert S.get_variable("w") is v
    with S.variable_scope("a"):                                        # entering by name again: reuse is the parent's
        with pytest.raises(ValueError):
            S.get_variable("w", [2])
    with S.variable_scope(sc):
        assert S.get_variable_scope().name == "a"
    assert v.var_name == "a/w"


def test_shim_session_feeds_intermediate_tensors_and_casts():
    S.reset_default_graph()
    p = S.placeholder(S.float32, [2, None], name="p")
    mid = p * 2.0
    st = S.LSTMStateTuple(mid + 1.0, mid - 1.0)
    out = st.c + st.h
    assert _run(out, {p: np.ones((2, 3))}).dtype == np.float32 and _run(out, {p: np.ones((2, 3))}).tolist() == [[4.0] * 3] * 2
    # feeding the tuple replaces both leaves; float64 input is cast to the tensors' dtype
    got = _run(out, {st: S.LSTMStateTuple(np.full((2, 3), 5.0), np.full((2, 3), 7.0))})
    assert got.dtype == np.float32 and got.tolist() == [[12.0] * 3] * 2
    with pytest.raises(ValueError):
        _run(out)                                                      # placeholder not fed


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/main/python/pointer-generator"),
                    reason="the reference tree exists only in the build container")
def test_generator_reproduces_the_committed_small_vectors(capsys):
    """Re-runs the reference's code through the shim and compares with the committed file, bit for bit."""
    mr = _load("make_reference_golden")
    saved = dict(sys.modules)
    try:
        R = mr.load_reference()
        _, _, rows, col = mr.small_case(R)
        got = mr.pack(col, 30, 8)
    finally:
        for k in list(sys.modules):
            if k not in saved and (k.split(".")[0] in ("tensorflow", "nltk", "pyrouge") or k in mr.REF_FILES):
                del sys.modules[k]
    r = np.load(os.path.join(HERE, "ref_small.npz"))
    for k in ("tokens", "lens", "logprobs", "avg", "step_ids", "step_logp", "n_steps"):
        assert np.array_equal(got[k], r[k]), k
    assert [x[2] for x in col.rows] == [str(s) for s in r["row_summary"]]
