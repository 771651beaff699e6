"""GPU parity: libpgen_cuda.so (through the C ABI) against the oracle on identical weights and inputs.

Bar (BASELINE.json north_star): token ids identical except at documented near-ties, log-probs within
1e-3 relative.  A "near-tie" is a beam-search step at which two candidates' summed log-probs differ by
less than NEAR_TIE relative, so that float32 rounding of a different (but equally valid) summation
order may swap them; such articles are counted and bounded, never silently accepted."""
import numpy as np
import pytest

import pgen_oracle as O
from conftest import small_flags
from textsummarization_on_flink_b200 import native, synth

pytestmark = pytest.mark.gpu

RTOL = 1e-3          # north_star: log-probs within 1e-3 relative
ATOL = 2e-5
NEAR_TIE = 2e-5


def _oracle_decode(W, hps, arts):
    out = []
    for ids, ext, n_oov in arts:
        tr = []
        toks, lps, avg = O.decode_ids(W, hps, ids, ext, n_oov, mode="dedup", trace=tr)
        out.append((toks, np.array(lps, np.float32), avg, tr))
    return out


def _has_near_tie(trace, upto):
    for step in trace[:upto + 1]:
        sc = np.array([x[1] for x in step], np.float64)
        sc = sc[np.isfinite(sc)]
        if len(sc) > 1:
            gap = np.abs(np.diff(sc)) / np.maximum(1e-12, np.abs(sc[:-1]))
            if gap.min() < NEAR_TIE:
                return True
    return False


def check_against_oracle(dec, W, hps, arts, max_near_tie_frac=0.1, decode=None):
    """Every article against the oracle.  Identical tokens and log-probs within RTOL -- or, at a documented near-tie,
    identical tokens and log-probs up to the step where the oracle's own candidates are within NEAR_TIE of each
    other; such articles are counted and bounded.  `decode`: alternative way of getting (tok, len, logp, avg)."""
    ids, ext, lens, noov = native.pack_articles(arts, dec.max_enc_steps)
    tok, ln, lp, avg = decode(ids, ext, lens, noov) if decode else dec.decode_ids(ids, ext, lens, noov)
    ref = _oracle_decode(W, hps, arts)
    near = 0
    for b, (rt, rl, ravg, tr) in enumerate(ref):
        got = tok[b, :ln[b]].tolist()
        if got != rt:
            first = next((i for i, (x, y) in enumerate(zip(got, rt)) if x != y), min(len(got), len(rt)))
            assert _has_near_tie(tr, min(first, len(tr) - 1)), \
                "article %d: tokens differ at %d with no near-tie\n got %s\n ref %s" % (b, first, got, rt)
            # the common prefix still has to carry the oracle's log-probs (the swap may have happened at an earlier
            # step than the one where the winning hypothesis shows it: allow the step before the difference)
            upto = max(0, first - 1)
            np.testing.assert_allclose(lp[b, :upto], rl[:upto], rtol=RTOL, atol=ATOL, err_msg="article %d prefix" % b)
            near += 1
            continue
        assert ln[b] == len(rt)
        np.testing.assert_allclose(lp[b, :ln[b]], rl, rtol=RTOL, atol=ATOL, err_msg="article %d log-probs" % b)
        assert abs(avg[b] - ravg) <= RTOL * abs(ravg) + ATOL
    assert near <= max(1, int(max_near_tie_frac * len(arts))), "%d of %d articles diverged at near-ties" % (near, len(arts))
    return near


@pytest.mark.parametrize("beam,coverage", [(4, 1), (4, 0), (1, 1), (3, 1), (8, 1)])
def test_small_model_matches_oracle(small_cfg, small_model_dir, beam, coverage):
    flags = small_flags(small_model_dir, small_cfg, beam=beam, coverage=coverage, extra="--articles_per_batch=16")
    hps = O.parse_hyperparams(flags)
    arts = synth.make_article_ids(24, vsize=small_cfg["V"], seed=100 + beam, max_len=24, oov_rate=0.15)
    arts.append((np.zeros(0, np.int32), np.zeros(0, np.int32), 0))          # zero-token article -> empty summary
    arts.append((np.array([7], np.int32), np.array([small_cfg["V"]], np.int32), 1))   # one OOV token only
    with native.Decoder(flags) as dec:
        check_against_oracle(dec, small_cfg["W"], hps, arts)


def test_small_model_no_pointer_gen(small_cfg, small_model_dir):
    flags = small_flags(small_model_dir, small_cfg, extra="--pointer_gen=0")
    hps = O.parse_hyperparams(flags)
    arts = synth.make_article_ids(10, vsize=small_cfg["V"], seed=5, max_len=24, oov_rate=0.1)
    with native.Decoder(flags) as dec:
        check_against_oracle(dec, small_cfg["W"], hps, arts)


def test_encoder_states_and_features(small_cfg, small_model_dir):
    flags = small_flags(small_model_dir, small_cfg, max_dec=1, min_dec=0)
    arts = synth.make_article_ids(9, vsize=small_cfg["V"], seed=3, max_len=24, oov_rate=0.1)
    W, H = small_cfg["W"], small_cfg["H"]
    with native.Decoder(flags) as dec:
        ids, ext, lens, noov = native.pack_articles(arts, dec.max_enc_steps)
        dec.stage_upload(ids, ext, lens, noov)
        dec.stage_run()
        dec.stage_sync()
        n, T, A = len(arts), dec.max_enc_steps, 2 * H
        es = dec.debug_fetch("enc_states", n * T * A).reshape(n, T, A)
        ef = dec.debug_fetch("enc_feat", n * T * A).reshape(n, T, A)
        c0 = dec.debug_fetch("dec_c0", n * H).reshape(n, H)
        h0 = dec.debug_fetch("dec_h0", n * H).reshape(n, H)
    for b, (a, _, _) in enumerate(arts):
        enc, (rc, rh) = O.run_encoder(W, a)
        L = len(a)
        np.testing.assert_allclose(es[b, :L], enc, atol=2e-6, rtol=1e-5)
        assert not es[b, L:].any()                                          # zero past enc_len (dynamic_rnn)
        np.testing.assert_allclose(ef[b, :L], O.encoder_features(W, enc[None])[0], atol=5e-6, rtol=1e-5)
        np.testing.assert_allclose(c0[b], rc, atol=2e-6, rtol=1e-5)
        np.testing.assert_allclose(h0[b], rh, atol=2e-6, rtol=1e-5)


def test_batch_composition_does_not_change_an_article(small_cfg, small_model_dir):
    """The reference decodes one article per batch (batcher.py:582); batching must not change results."""
    flags = small_flags(small_model_dir, small_cfg, extra="--articles_per_batch=8")
    arts = synth.make_article_ids(8, vsize=small_cfg["V"], seed=9, max_len=24, oov_rate=0.15)
    with native.Decoder(flags) as dec:
        ids, ext, lens, noov = native.pack_articles(arts, dec.max_enc_steps)
        tok, ln, lp, avg = dec.decode_ids(ids, ext, lens, noov)
        for b in range(len(arts)):
            t1, l1, p1, a1 = dec.decode_ids(ids[b:b + 1], ext[b:b + 1], lens[b:b + 1], noov[b:b + 1])
            assert l1[0] == ln[b] and np.array_equal(t1[0], tok[b])
            assert np.array_equal(p1[0, :l1[0]], lp[b, :ln[b]]) and a1[0] == avg[b]   # bit-identical


def test_full_size_w1_matches_oracle(model_dir):
    """Config 1 of BASELINE.json: beam 4, 400/100 steps, coverage, vocab 50k, hidden 256, emb 128."""
    flags = synth.hyperparams_string(model_dir, extra="--articles_per_batch=8")
    hps = O.parse_hyperparams(flags)
    from textsummarization_on_flink_b200 import weights_io
    W = weights_io.load_weights(model_dir["weights"])
    arts = synth.make_article_ids(6, seed=77)
    arts.append(synth.make_article_ids(1, seed=5, max_len=3)[0])
    with native.Decoder(flags) as dec:
        check_against_oracle(dec, W, hps, arts)
        st = dec.stats()
        assert st["kernel_launches"] > 0 and st["batches"] >= 1


def test_rows_submit_poll_matches_oracle_rows(small_cfg, small_model_dir):
    """open -> N x submit/poll -> flush -> close, the operator's call pattern."""
    flags = small_flags(small_model_dir, small_cfg, extra="--articles_per_batch=4 --flush_ms=-1")
    hps = O.parse_hyperparams(flags)
    vocab = O.Vocab(small_model_dir["vocab_path"], small_cfg["V"])
    arts = synth.make_articles(10, vsize=small_cfg["V"], seed=21, max_len=24, oov_rate=0.2)
    arts += [b"article 0.", b"", b"   "]                                     # TensorFlowTest.java:209; empty rows
    got = []
    with native.Decoder(flags) as dec:
        for i, a in enumerate(arts):
            dec.submit(b"uuid-%d" % i, a, b"reference %d." % i)
            got += dec.poll()
        # decoding is asynchronous: poll() hands back only batches that have finished, never a partial one
        assert len(got) % 4 == 0 and len(got) <= (len(arts) // 4) * 4
        got += dec.poll(flush=True)
        st = dec.stats()
    assert st["rows_in"] == st["rows_out"] == len(arts)
    assert [r[0] for r in got] == [b"uuid-%d" % i for i in range(len(arts))]  # submission order
    for i, (uuid, article, summary, reference) in enumerate(got):
        ref = O.decode_row(small_cfg["W"], vocab, hps, b"uuid-%d" % i, arts[i], b"reference %d." % i, mode="dedup")
        assert (uuid, article, reference) == (ref[0], ref[1], ref[3])        # echoed unchanged
        assert summary == ref[2], "row %d" % i


@pytest.mark.parametrize("encoder", ["persistent", "steps"])
def test_full_size_encoder_states(model_dir, encoder):
    """Config 2 of BASELINE.json (encoder only) at full size: the persistent cluster kernel (hidden 256) and the
    per-time-step kernel against the oracle, ragged lengths included."""
    from textsummarization_on_flink_b200 import weights_io
    flags = synth.hyperparams_string(model_dir, max_dec_steps=1, min_dec_steps=0, extra="--articles_per_batch=16")
    W = weights_io.load_weights(model_dir["weights"])
    arts = synth.make_article_ids(9, seed=31) + synth.make_article_ids(2, seed=32, max_len=5)
    arts.append((np.zeros(0, np.int32), np.zeros(0, np.int32), 0))
    H = 256
    with native.Decoder(flags) as dec:
        dec.set_option("encoder=" + encoder)
        ids, ext, lens, noov = native.pack_articles(arts, dec.max_enc_steps)
        dec.stage_upload(ids, ext, lens, noov)
        dec.stage_run()
        dec.stage_sync()
        n, T, A = len(arts), dec.max_enc_steps, 2 * H
        es = dec.debug_fetch("enc_states", n * T * A).reshape(n, T, A)
        ef = dec.debug_fetch("enc_feat", n * T * A).reshape(n, T, A)
        c0 = dec.debug_fetch("dec_c0", n * H).reshape(n, H)
        h0 = dec.debug_fetch("dec_h0", n * H).reshape(n, H)
    for b, (a, _, _) in enumerate(arts):
        L = len(a)
        assert not es[b, L:].any()
        if L == 0:
            continue
        enc, (rc, rh) = O.run_encoder(W, a)
        np.testing.assert_allclose(es[b, :L], enc, atol=2e-5, rtol=1e-4)
        np.testing.assert_allclose(ef[b, :L], O.encoder_features(W, enc[None])[0], atol=1e-4, rtol=1e-4)
        np.testing.assert_allclose(c0[b], rc, atol=2e-5, rtol=1e-4)
        np.testing.assert_allclose(h0[b], rh, atol=2e-5, rtol=1e-4)


def test_config4_beam8_200_oovs(model_dir):
    """Config 4 of BASELINE.json: beam 8, max_dec_steps 120, exactly 200 distinct OOV words per article
    (extended vocabulary 50 200, top-16)."""
    from textsummarization_on_flink_b200 import weights_io
    flags = synth.hyperparams_string(model_dir, beam_size=8, max_dec_steps=120, extra="--articles_per_batch=4")
    hps = O.parse_hyperparams(flags)
    W = weights_io.load_weights(model_dir["weights"])
    arts = synth.make_article_ids(3, seed=404, fixed_oovs=200, full_len_only=True)
    assert all(k == 200 for _, _, k in arts)
    with native.Decoder(flags) as dec:
        check_against_oracle(dec, W, hps, arts, max_near_tie_frac=0.34)
    # the same articles inside a full device batch of 64 (4 row blocks of 128 in the tensor-core kernel)
    flags64 = synth.hyperparams_string(model_dir, beam_size=8, max_dec_steps=120, extra="--articles_per_batch=64 --slots=1")
    more = synth.make_article_ids(61, seed=405, fixed_oovs=200)
    with native.Decoder(flags64) as dec:
        ids, ext, lens, noov = native.pack_articles(arts + more, dec.max_enc_steps)
        tok, ln, lp, avg = dec.decode_ids(ids, ext, lens, noov)
        with native.Decoder(flags) as small:
            t2, l2, p2, a2 = small.decode_ids(ids[:3], ext[:3], lens[:3], noov[:3])
        for b in range(3):
            assert ln[b] == l2[b] and np.array_equal(tok[b], t2[b])
            np.testing.assert_allclose(lp[b, :ln[b]], p2[b, :l2[b]], rtol=1e-4, atol=1e-5)


def test_w0_reference_scale_weights_logprobs(tmp_path_factory):
    """W0 = the reference's own initialisers (trunc-normal 1e-4, LSTM U(-0.02, 0.02)): the 50k-way distribution is
    almost flat (logits within ~1e-3 of each other), so nearly every beam step holds candidates closer than NEAR_TIE
    and WHICH of them wins is rounding (SURVEY.md hard parts).  Same accounting as everywhere else: tokens identical, or
    identical up to a step with a near-tie in the oracle's own candidates, log-probs within 1e-3 on the common
    prefix; the first emitted token and the average log-prob are well defined and must always agree."""
    root = str(tmp_path_factory.mktemp("pgen_w0"))
    paths = synth.make_model_dir(root, "W0")
    from textsummarization_on_flink_b200 import weights_io
    W = weights_io.load_weights(paths["weights"])
    flags = synth.hyperparams_string(paths, max_dec_steps=40, min_dec_steps=10, extra="--articles_per_batch=8")
    hps = O.parse_hyperparams(flags)
    arts = synth.make_article_ids(6, seed=91)
    with native.Decoder(flags) as dec:
        got = {}

        def decode(*packed):
            got["out"] = dec.decode_ids(*packed)
            return got["out"]

        check_against_oracle(dec, W, hps, arts, max_near_tie_frac=1.0, decode=decode)
        tok, ln, lp, avg = got["out"]
    for b, (a, e, k) in enumerate(arts):
        rt, rl, ravg = O.decode_ids(W, hps, a, e, k, mode="dedup")
        assert ln[b] == len(rt)
        assert abs(lp[b, 0] - rl[0]) <= RTOL * abs(rl[0]) + ATOL
        assert abs(avg[b] - ravg) <= RTOL * abs(ravg) + 1e-4


def test_edge_rows_through_submit(small_cfg, small_model_dir):
    """Truncation at max_enc_steps (batcher.py:52-54), an all-OOV article, whitespace-only and empty rows, empty
    uuid / reference: rows come back in order and equal the oracle's rows."""
    flags = small_flags(small_model_dir, small_cfg, extra="--articles_per_batch=3 --flush_ms=-1")
    hps = O.parse_hyperparams(flags)
    vocab = O.Vocab(small_model_dir["vocab_path"], small_cfg["V"])
    long_article = b" ".join(b"w%d" % (4 + i % 200) for i in range(500)) + b" tail1 tail2"     # 502 words > 24
    rows = [(b"a", long_article, b"r a"), (b"", b"zzz yyy zzz xxx yyy", b""), (b"c", b" \t\n ", b"r c"),
            (b"d", b"", b""), (b"e", b". . . , ,", b"r e"), (b"f", b"w7", b"r f"), (b"g", b"w8 " * 24, b"r g")]
    with native.Decoder(flags) as dec:
        got = []
        for u, a, r in rows:
            dec.submit(u, a, r)
            got += dec.poll()
        got += dec.poll(flush=True)
        assert dec.stats()["tokens_in"] == sum(min(24, len(a.split())) for _, a, _ in rows)
    assert len(got) == len(rows)
    for (u, a, r), g in zip(rows, got):
        assert g == O.decode_row(small_cfg["W"], vocab, hps, u, a, r, mode="dedup")
    assert got[2][2] == b"" and got[3][2] == b""                   # zero-token articles: empty summary


def test_early_stop_polling_does_not_change_results(small_cfg, small_model_dir):
    flags = small_flags(small_model_dir, small_cfg, extra="--articles_per_batch=8")
    arts = synth.make_article_ids(8, vsize=small_cfg["V"], seed=13, max_len=24, oov_rate=0.1)
    with native.Decoder(flags) as dec:
        ref = dec.decode_ids(*native.pack_articles(arts, dec.max_enc_steps))
        dec.set_option("check_every=2")
        got = dec.decode_ids(*native.pack_articles(arts, dec.max_enc_steps))
    for x, y in zip(ref, got):
        assert np.array_equal(x, y)


def test_word_table_and_graph_replay_do_not_change_results(model_dir):
    """The per-word input-projection table (--encoder_table) and the CUDA-graph replay of the decode steps are
    pure re-arrangements: tokens and log-probs are bit-identical with either switched off."""
    arts = synth.make_article_ids(6, seed=31)
    outs = []
    for extra, opts in (("--articles_per_batch=8", []), ("--articles_per_batch=8 --encoder_table=0", []),
                        ("--articles_per_batch=8", ["graph=0"]), ("--articles_per_batch=8", ["graph_steps=4"])):
        with native.Decoder(synth.hyperparams_string(model_dir, extra=extra)) as dec:
            for o in opts:
                dec.set_option(o)
            packed = native.pack_articles(arts, dec.max_enc_steps)
            dec.decode_ids(*packed)                      # first batch: plain launches, then the capture
            outs.append(dec.decode_ids(*packed))         # second batch: replay
    for got in outs[1:]:
        for x, y in zip(outs[0], got):
            assert np.array_equal(x, y)


def test_features_on_tensor_cores_match_fp32(model_dir):
    """W_h.h_i by the 3xTF32 tcgen05 GEMM against the fp32 SIMT GEMM on the same encoder states, and against a float64
    product of those states.  The split operands are exact to 2^-22, but the tensor core's fp32 accumulator truncates, so
    over K = 512 the result is ~1e-5 absolute from the exact product on O(1) values (the fp32 FMA chain: ~1e-6); the
    bound below is what the decode parity (log-probs within 1e-3) is run with."""
    from textsummarization_on_flink_b200 import weights_io
    flags = synth.hyperparams_string(model_dir, max_dec_steps=1, min_dec_steps=0, extra="--articles_per_batch=16")
    W = weights_io.load_weights(model_dir["weights"])
    Wh = W["seq2seq/decoder/attention_decoder/W_h"].reshape(512, 512).astype(np.float64)
    arts = synth.make_article_ids(16, seed=41)
    got = {}
    with native.Decoder(flags) as dec:
        packed = native.pack_articles(arts, dec.max_enc_steps)
        n, T, A = len(arts), dec.max_enc_steps, 512
        for mode in ("tc", "simt"):
            dec.set_option("features=" + mode)
            dec.stage_upload(*packed)
            dec.stage_run()
            dec.stage_sync()
            got[mode] = dec.debug_fetch("enc_feat", n * T * A).reshape(n, T, A)
        es = dec.debug_fetch("enc_states", n * T * A).reshape(n, T, A)
    err = {}
    for b, (a, _, _) in enumerate(arts):
        L = len(a)
        exact = es[b, :L].astype(np.float64) @ Wh
        for mode in got:
            err[mode] = max(err.get(mode, 0.0), float(np.abs(got[mode][b, :L] - exact).max()))
        np.testing.assert_allclose(got["tc"][b, :L], got["simt"][b, :L], atol=2e-5, rtol=1e-4)
    assert err["simt"] <= 5e-6 and err["tc"] <= 2e-5, err
