"""The top-2K of a decoder row must be tf.nn.top_k over the whole extended vocabulary (model.py:284-285).

The tensor-core pass only *selects* candidate groups from tf32 scores; vocab_final_kernel then checks every row
against an error bound and widens the candidate set (or scores the whole vocabulary in fp32) when an outside word
could reach the 2K-th value.  These tests drive that guard through the C ABI:

* the tensor-core path against the dense all-fp32 path of the library (`vocab=simt`) on the beam traces of 64
  full-size articles, with the reference's own initialisers (W0, near-flat distributions) and with W1;
* the dense all-fp32 path against the oracle;
* forced `widen` / `dense` modes against `auto`;
* plateaus of exactly tied logits (zero projection columns + equal bias), where the tf32 group maxima cannot tell
  the groups apart: the result must follow tf.nn.top_k's tie rule (lower id first), as the oracle does."""
import os

import numpy as np
import pytest

import pgen_oracle as O
from conftest import small_flags
from test_gpu_parity import check_against_oracle, NEAR_TIE
from textsummarization_on_flink_b200 import native, synth, weights_io

pytestmark = pytest.mark.gpu

LP_TOL = 5e-4       # see compare_traces
SUM_TOL = 1e-3      # two runs' k-th best summed log-probs at a step where their beams first differ


def _traces(dec, packed, n):
    dec.stage_upload(*packed)
    dec.stage_run()
    dec.stage_sync()
    S, K = dec.max_dec_steps, dec.beam_size
    tok = dec.debug_fetch("trace_tokens", n * S * K, np.int32).reshape(n, S, K).copy()
    par = dec.debug_fetch("trace_parents", n * S * K, np.int32).reshape(n, S, K).copy()
    lp = dec.debug_fetch("trace_logp", n * S * K).reshape(n, S, K).copy()
    steps = dec.debug_fetch("steps_done", n, np.int32).copy()
    out = dec.stage_download()
    return tok, par, lp, steps, out


def _beam_sums(par, lp, upto):
    """Summed log-prob of every live hypothesis after each step < upto (beam_search.py:72-74)."""
    K = par.shape[1]
    sums = np.zeros(K, np.float64)
    hist = []
    for s in range(upto):
        sums = sums[par[s]] + lp[s].astype(np.float64)
        hist.append(sums.copy())
    return hist


def compare_traces(a, b, n, max_diverged):
    """Two runs of the same articles: identical beams, or a first difference at a step where the two runs' k-th best
    summed log-probs agree within SUM_TOL (a swap among near-equal candidates).  Log-probs within LP_TOL up to there.

    LP_TOL is not fp32 rounding: the two paths share every selected value, but the tensor-core path's softmax
    denominator keeps the tf32 score of every word it did NOT re-score (the re-scored ones are swapped for their exact
    value).  With W1 a few hundred words just below the candidates carry ~1e-4 of relative error into that sum, which
    shifts every vocabulary term of the row by the same factor (measured: <= 2.6e-4 absolute on log-probs)."""
    tok_a, par_a, lp_a, steps_a, _ = a
    tok_b, par_b, lp_b, steps_b, _ = b
    diverged = 0
    for i in range(n):
        ns = int(min(steps_a[i], steps_b[i]))
        first = None
        for s in range(ns):
            if not (np.array_equal(tok_a[i, s], tok_b[i, s]) and np.array_equal(par_a[i, s], par_b[i, s])):
                first = s
                break
        upto = ns if first is None else first
        np.testing.assert_allclose(lp_a[i, :upto], lp_b[i, :upto], rtol=LP_TOL, atol=LP_TOL, err_msg="article %d" % i)
        if first is None:
            assert steps_a[i] == steps_b[i]
            continue
        sa = _beam_sums(par_a[i], lp_a[i], first + 1)[-1]
        sb = _beam_sums(par_b[i], lp_b[i], first + 1)[-1]
        # same number of survivors with pairwise near-equal sums: a swap among near-equal candidates
        np.testing.assert_allclose(np.sort(sa), np.sort(sb), rtol=SUM_TOL, atol=SUM_TOL,
                                   err_msg="article %d diverges at step %d without a near-tie" % (i, first))
        diverged += 1
    assert diverged <= max_diverged, "%d of %d articles diverged" % (diverged, n)
    return diverged


@pytest.mark.parametrize("kind", ["W1", "W0"])
def test_tensor_core_path_matches_dense_fp32_path(tmp_path_factory, model_dir, kind):
    paths = model_dir if kind == "W1" else synth.make_model_dir(str(tmp_path_factory.mktemp("pgen_w0g")), "W0")
    n = 64
    arts = synth.make_article_ids(n, seed=515)
    flags = synth.hyperparams_string(paths, max_dec_steps=60, min_dec_steps=20, extra="--articles_per_batch=64 --slots=1")
    with native.Decoder(flags) as dec:
        packed = native.pack_articles(arts, dec.max_enc_steps)
        tc = _traces(dec, packed, n)
        st = dec.stats()
        dec.set_option("vocab=simt")
        simt = _traces(dec, packed, n)
    # W0: ~50k words within 1e-3 of each other -- swaps at rounding level are expected, undetected misses are not
    div = compare_traces(tc, simt, n, max_diverged=n if kind == "W0" else 6)
    print("%s: %d/%d articles swap a near-tie; widened rows %d, dense rows %d of %d row-steps"
          % (kind, div, n, st["selection_widened"], st["selection_fallbacks"], 4 * int(tc[3].sum())))


def test_dense_fp32_path_matches_oracle(model_dir):
    flags = synth.hyperparams_string(model_dir, extra="--articles_per_batch=8")
    hps = O.parse_hyperparams(flags)
    W = weights_io.load_weights(model_dir["weights"])
    arts = synth.make_article_ids(5, seed=78)
    with native.Decoder(flags) as dec:
        dec.set_option("vocab=simt")
        check_against_oracle(dec, W, hps, arts)


@pytest.mark.parametrize("mode", ["widen", "dense"])
def test_forced_guard_modes_change_nothing(model_dir, mode):
    arts = synth.make_article_ids(8, seed=616)
    flags = synth.hyperparams_string(model_dir, max_dec_steps=40, min_dec_steps=15, extra="--articles_per_batch=8 --slots=1")
    with native.Decoder(flags) as dec:
        packed = native.pack_articles(arts, dec.max_enc_steps)
        auto = _traces(dec, packed, len(arts))
        dec.set_option("vocab_guard=" + mode)
        forced = _traces(dec, packed, len(arts))
        st = dec.stats()
    assert st["selection_widened"] + st["selection_fallbacks"] > 0
    if mode == "dense":
        assert st["selection_fallbacks"] >= int(forced[3].sum())         # every row of every step went dense
    compare_traces(auto, forced, len(arts), max_diverged=1)
    for x, y in zip(auto[4][:2], forced[4][:2]):
        assert np.array_equal(x, y)                                      # winning tokens and lengths


def _plateau_model(tmp_path_factory, V, H, E, lo, hi, bias, seed):
    """W1-style weights whose vocabulary words [lo, hi) all score exactly `bias` (zero column + equal bias)."""
    W = synth.make_weights("W1", vsize=V, hidden=H, emb=E, seed=seed)
    W["seq2seq/output_projection/w"][:, lo:hi] = 0.0
    W["seq2seq/output_projection/v"][lo:hi] = np.float32(bias)
    root = str(tmp_path_factory.mktemp("pgen_plateau"))
    train = os.path.join(root, "log", "plateau", "train")
    os.makedirs(train)
    weights_io.save_weights(os.path.join(train, weights_io.DEFAULT_NAME), W)
    synth.write_vocab_file(os.path.join(root, "vocab"), V)
    return W, {"vocab_path": os.path.join(root, "vocab"), "log_root": os.path.join(root, "log"), "exp_name": "plateau"}


@pytest.mark.parametrize("V,lo,hi,expect", [(300, 40, 290, "widened"), (2400, 100, 2300, "dense")])
def test_tied_plateau_follows_top_k_tie_rule(tmp_path_factory, V, lo, hi, expect):
    """Hundreds of words with EXACTLY equal scores above everything else: the tf32 group maxima are all equal, so
    which groups get selected is arbitrary; tf.nn.top_k returns the lowest ids.  The guard must notice and the
    tokens must equal the oracle's."""
    H, E = 32, 16
    W, paths = _plateau_model(tmp_path_factory, V, H, E, lo, hi, bias=30.0, seed=3)
    cfg = dict(V=V, H=H, E=E)
    flags = small_flags(paths, cfg, beam=4, coverage=1, max_dec=12, min_dec=4, extra="--articles_per_batch=8")
    hps = O.parse_hyperparams(flags)
    # short articles: at most 3 words carry copy mass, so the top 8 must be filled from the tied plateau
    arts = synth.make_article_ids(8, vsize=V, seed=44, max_len=3, oov_rate=0.1)
    with native.Decoder(flags) as dec:
        check_against_oracle(dec, W, hps, arts, max_near_tie_frac=0.0)
        st = dec.stats()
    assert st["selection_" + ("widened" if expect == "widened" else "fallbacks")] > 0, st
