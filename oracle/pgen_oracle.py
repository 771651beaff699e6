"""CPU oracle for the pointer-generator beam-search decode path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and there only as the checker or the timed
CPU baseline -- never as the thing shipped.  The product path is ``libpgen_cuda.so``.

PARITY STATUS.  The reference's arithmetic lives in TensorFlow 1.x ("tensorflow
1.11.0", /root/reference/README.md:87), which is not vendored under /root/reference and
is not installable here (Python 2.7 code, no network); the reference's own tests hold no
golden vectors or assertions.  This file is therefore a
*restatement* of the reference algorithm from its call sites, with the TF1 op semantics
written out from their published definitions.  Each function cites the reference
file:line it follows (paths relative to /root/reference/src/main/python/pointer-generator/).
What pins it:
* outputs of the reference's OWN sources run here: model.py, attention_decoder.py,
  beam_search.py, decode.py, batcher.py, data.py executed unmodified on oracle/tf1_shim.py
  (a NumPy restatement of the TF 1.11 ops they call) by
  tests/golden/make_reference_golden.py; the vectors are committed (tests/golden/ref_*.npz)
  and tests/test_reference_golden.py holds this file to them step by step and row by row
  (identical token ids; log-probs within 2e-5 on the small model, 5e-4 at full size);
* tests/test_oracle_*.py: every op against an independent formulation (torch.nn.LSTM,
  torch.softmax, index_add_, torch.topk, a float64 whole-decode), and the two oracle
  modes ("literal" and "dedup") against each other.
Still unpinned: TensorFlow's own kernels (the shim restates their documented semantics;
no TensorFlow binary has run on these inputs).

Arithmetic: float32 throughout (NumPy), beam scores accumulated in float64 exactly as the
reference's Python ``sum`` over float32 scalars does (beam_search.py:72-79).
"""
from __future__ import annotations

import dataclasses
import math
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

F32 = np.float32

# --------------------------------------------------------------------------------------
# Text path (data.py, batcher.py, decode.py)
# --------------------------------------------------------------------------------------

SENTENCE_START = "<s>"          # data.py:26
SENTENCE_END = "</s>"           # data.py:27
PAD_TOKEN = "[PAD]"             # data.py:29
UNKNOWN_TOKEN = "[UNK]"         # data.py:30
START_DECODING = "[START]"      # data.py:31
STOP_DECODING = "[STOP]"        # data.py:32
UNK_ID, PAD_ID, START_ID, STOP_ID = 0, 1, 2, 3   # data.py:50-54


class Vocab:
    """data.py:37-91.  ids 0..3 are the specials, then file order, capped at max_size."""

    def __init__(self, vocab_file: str, max_size: int):
        self._word_to_id: Dict[bytes, int] = {}
        self._id_to_word: Dict[int, bytes] = {}
        self._count = 0
        self.warnings: List[str] = []
        for w in (UNKNOWN_TOKEN, PAD_TOKEN, START_DECODING, STOP_DECODING):   # data.py:50-54
            self._add(w.encode())
        forbidden = {s.encode() for s in (SENTENCE_START, SENTENCE_END, UNKNOWN_TOKEN,
                                           PAD_TOKEN, START_DECODING, STOP_DECODING)}
        with open(vocab_file, "rb") as f:                                     # data.py:57-73
            for line in f:
                pieces = line.split()
                if len(pieces) != 2:                                          # data.py:59-62
                    self.warnings.append("incorrectly formatted line: %r" % line)
                    continue
                w = pieces[0]
                if w in forbidden:                                            # data.py:64-65
                    raise ValueError("special token %r must not be in the vocab file" % w)
                if w in self._word_to_id:                                     # data.py:66-67
                    raise ValueError("Duplicated word in vocabulary file: %r" % w)
                self._add(w)
                if max_size != 0 and self._count >= max_size:                 # data.py:71-73
                    break

    def _add(self, w: bytes) -> None:
        self._word_to_id[w] = self._count
        self._id_to_word[self._count] = w
        self._count += 1

    def word2id(self, word: bytes) -> int:            # data.py:77-81
        return self._word_to_id.get(word, UNK_ID)

    def id2word(self, word_id: int) -> bytes:         # data.py:83-87
        if word_id not in self._id_to_word:
            raise ValueError("Id not found in vocab: %d" % word_id)
        return self._id_to_word[word_id]

    def size(self) -> int:                            # data.py:89-91
        return self._count


def article2ids(article_words: Sequence[bytes], vocab: Vocab) -> Tuple[List[int], List[bytes]]:
    """data.py:144-168.  j-th distinct OOV (first-appearance order) -> vocab.size()+j."""
    ids: List[int] = []
    oovs: List[bytes] = []
    for w in article_words:
        i = vocab.word2id(w)
        if i == UNK_ID:
            if w not in oovs:
                oovs.append(w)
            ids.append(vocab.size() + oovs.index(w))
        else:
            ids.append(i)
    return ids, oovs


def outputids2words(id_list: Sequence[int], vocab: Vocab, article_oovs: Sequence[bytes]) -> List[bytes]:
    """data.py:196-219."""
    words = []
    for i in id_list:
        if i < vocab.size():
            words.append(vocab.id2word(i))
        else:
            idx = i - vocab.size()
            if idx >= len(article_oovs):
                raise ValueError("model produced word ID %d (article OOV %d) but the example has %d OOVs"
                                 % (i, idx, len(article_oovs)))
            words.append(article_oovs[idx])
    return words


@dataclasses.dataclass
class Example:
    """batcher.py:36-81 (encoder side only; decode never reads the abstract tensors)."""
    enc_len: int
    enc_input: List[int]
    enc_input_extend_vocab: List[int]
    article_oovs: List[bytes]


def make_example(article: bytes, vocab: Vocab, max_enc_steps: int) -> Example:
    words = article.split()                                  # batcher.py:51 (bytes.split(): ASCII whitespace)
    if len(words) > max_enc_steps:                           # batcher.py:52-54
        words = words[:max_enc_steps]
    enc_input = [vocab.word2id(w) for w in words]            # batcher.py:56
    ext, oovs = article2ids(words, vocab)                    # batcher.py:70
    return Example(len(words), enc_input, ext, oovs)


def cut_at_stop(decoded_words: List[bytes]) -> List[bytes]:
    """decode.py:113-118."""
    stop = STOP_DECODING.encode()
    if stop in decoded_words:
        return decoded_words[:decoded_words.index(stop)]
    return decoded_words


def format_summary(decoded_words: Sequence[bytes]) -> bytes:
    """decode.py:159-185: a sentence ends after each "." token; sentences joined by two spaces;
    make_html_safe (decode.py:261-265) discards the result of str.replace, i.e. is a no-op."""
    words = list(decoded_words)
    sents = []
    while words:
        try:
            k = words.index(b".")
        except ValueError:
            k = len(words)
        sents.append(b" ".join(words[:k + 1]))
        words = words[k + 1:]
    return b"  ".join(sents)


# --------------------------------------------------------------------------------------
# Hyper-parameter string (run_summarization.py:48-88, 312-313, 418-420)
# --------------------------------------------------------------------------------------

_FLAG_DEFAULTS = {            # name -> (type, default)   run_summarization.py:48-88
    "data_path": (str, ""), "vocab_path": (str, ""), "mode": (str, "train"),
    "num_steps": (int, 0), "single_pass": (bool, False), "inference": (bool, False),
    "log_root": (str, ""), "exp_name": (str, ""), "hidden_dim": (int, 256), "emb_dim": (int, 128),
    "batch_size": (int, 16), "max_enc_steps": (int, 400), "max_dec_steps": (int, 100),
    "beam_size": (int, 4), "min_dec_steps": (int, 35), "vocab_size": (int, 50000),
    "lr": (float, 0.15), "adagrad_init_acc": (float, 0.1), "rand_unif_init_mag": (float, 0.02),
    "trunc_norm_init_std": (float, 1e-4), "max_grad_norm": (float, 2.0),
    "pointer_gen": (bool, True), "coverage": (bool, False), "cov_loss_wt": (float, 1.0),
    "convert_to_coverage_model": (bool, False), "restore_best_model": (bool, False),
    "debug": (bool, False),
}


def _parse_bool(v: str) -> bool:
    lv = v.lower()
    if lv in ("1", "true", "t", "yes", "y"):
        return True
    if lv in ("0", "false", "f", "no", "n"):
        return False
    raise ValueError("bad boolean flag value %r" % v)


def parse_hyperparams(s: str) -> Dict[str, object]:
    """tf.app.flags.FLAGS(s.split(' '), known_only=True) (run_summarization.py:420): token 0 is a
    dummy program name (TensorFlowTest.java:41); unknown flags are ignored; in decode mode
    batch_size is overwritten by beam_size (run_summarization.py:312-313)."""
    flags = {k: d for k, (_, d) in _FLAG_DEFAULTS.items()}
    toks = s.split(" ")
    for tok in toks[1:]:
        if not tok.startswith("--"):
            continue
        body = tok[2:]
        if "=" in body:
            name, val = body.split("=", 1)
        else:
            name, val = body, None
        if name.startswith("no") and name[2:] in _FLAG_DEFAULTS and _FLAG_DEFAULTS[name[2:]][0] is bool and val is None:
            flags[name[2:]] = False
            continue
        if name not in _FLAG_DEFAULTS:
            continue
        typ = _FLAG_DEFAULTS[name][0]
        if typ is bool:
            flags[name] = True if val is None else _parse_bool(val)
        elif val is None:
            raise ValueError("flag --%s needs a value" % name)
        else:
            flags[name] = typ(val)
    if flags["mode"] == "decode":
        flags["batch_size"] = flags["beam_size"]
    return flags


# --------------------------------------------------------------------------------------
# Numerics (model.py, attention_decoder.py); TF1 op semantics per SURVEY.md section 8c
# --------------------------------------------------------------------------------------

P = "seq2seq/"
N_EMB = P + "embedding/embedding"
N_ENC = P + "encoder/bidirectional_rnn/%s/lstm_cell/%s"
N_RED = P + "reduce_final_st/%s"
D = P + "decoder/attention_decoder/"
N_WH, N_V, N_WC = D + "W_h", D + "v", D + "coverage/w_c"
N_ATT_M, N_ATT_B = D + "Attention/Linear/Matrix", D + "Attention/Linear/Bias"
N_X_M, N_X_B = D + "Linear/Matrix", D + "Linear/Bias"
N_DEC_K, N_DEC_B = D + "lstm_cell/kernel", D + "lstm_cell/bias"
N_PG_M, N_PG_B = D + "calculate_pgen/Linear/Matrix", D + "calculate_pgen/Linear/Bias"
N_OUT_M, N_OUT_B = D + "AttnOutputProjection/Linear/Matrix", D + "AttnOutputProjection/Linear/Bias"
N_PROJ_W, N_PROJ_V = P + "output_projection/w", P + "output_projection/v"


def variable_shapes(vsize: int, hidden: int = 256, emb: int = 128) -> Dict[str, Tuple[int, ...]]:
    """Variable inventory of the decode graph."""
    H, E, A = hidden, emb, 2 * hidden
    sh = {N_EMB: (vsize, E)}
    for d in ("fw", "bw"):
        sh[N_ENC % (d, "kernel")] = (E + H, 4 * H)
        sh[N_ENC % (d, "bias")] = (4 * H,)
    sh[N_RED % "w_reduce_c"] = (A, H)
    sh[N_RED % "w_reduce_h"] = (A, H)
    sh[N_RED % "bias_reduce_c"] = (H,)
    sh[N_RED % "bias_reduce_h"] = (H,)
    sh[N_WH] = (1, 1, A, A)
    sh[N_V] = (A,)
    sh[N_WC] = (1, 1, 1, A)
    sh[N_ATT_M], sh[N_ATT_B] = (2 * H, A), (A,)
    sh[N_X_M], sh[N_X_B] = (E + A, E), (E,)
    sh[N_DEC_K], sh[N_DEC_B] = (E + H, 4 * H), (4 * H,)
    sh[N_PG_M], sh[N_PG_B] = (A + H + H + E, 1), (1,)
    sh[N_OUT_M], sh[N_OUT_B] = (H + A, H), (H,)
    sh[N_PROJ_W], sh[N_PROJ_V] = (H, vsize), (vsize,)
    return sh


def sigmoid(x: np.ndarray) -> np.ndarray:
    return (F32(1.0) / (F32(1.0) + np.exp(-x, dtype=F32))).astype(F32)


def lstm_cell(x, c, h, kernel, bias):
    """tf.contrib.rnn.LSTMCell (model.py:90-91,138): z=[x;h].kernel+bias; split order i,j,f,o;
    c' = sigmoid(f+1)*c + sigmoid(i)*tanh(j); h' = sigmoid(o)*tanh(c').  forget_bias=1.0."""
    z = np.concatenate([x, h], axis=-1).astype(F32) @ kernel + bias
    i, j, f, o = np.split(z, 4, axis=-1)
    c2 = sigmoid(f + F32(1.0)) * c + sigmoid(i) * np.tanh(j, dtype=F32)
    h2 = sigmoid(o) * np.tanh(c2, dtype=F32)
    return c2.astype(F32), h2.astype(F32)


def softmax(x: np.ndarray) -> np.ndarray:
    """nn_ops.softmax over the last axis (attention_decoder.py:98, model.py:238)."""
    m = x.max(axis=-1, keepdims=True)
    e = np.exp(x - m, dtype=F32)
    return (e / e.sum(axis=-1, keepdims=True, dtype=F32)).astype(F32)


def top_k(x: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
    """tf.nn.top_k (model.py:284): descending values; among equal values the lower index first."""
    n = x.shape[0]
    k = min(k, n)
    if n > 4 * k:
        kth = np.partition(x, n - k)[n - k]
        cand = np.nonzero(x >= kth)[0]
    else:
        cand = np.arange(n)
    order = np.lexsort((cand, -x[cand].astype(np.float64)))[:k]
    idx = cand[order]
    return x[idx], idx.astype(np.int32)


def run_encoder(W: Dict[str, np.ndarray], enc_ids: np.ndarray):
    """model.py:210-221,76-121,347-364 for ONE article of length T (no padding: in decode mode
    the batch is the article repeated, batcher.py:582, so enc_len == T for every row).
    Returns enc_states [T,2H], (c0,h0) [H]."""
    T = len(enc_ids)
    emb = W[N_EMB][np.asarray(enc_ids, dtype=np.int64)]          # model.py:213 (ids already OOV->UNK)
    H = W[N_RED % "bias_reduce_c"].shape[0]
    out = np.zeros((T, 2 * H), dtype=F32)
    finals = {}
    for d, order in (("fw", range(T)), ("bw", range(T - 1, -1, -1))):   # bw: reverse_sequence over own length
        k, b = W[N_ENC % (d, "kernel")], W[N_ENC % (d, "bias")]
        c = np.zeros((H,), dtype=F32)
        h = np.zeros((H,), dtype=F32)
        off = 0 if d == "fw" else H
        for t in order:
            c, h = lstm_cell(emb[t], c, h, k, b)
            out[t, off:off + H] = h                              # model.py:93 concat [fw;bw]
        finals[d] = (c, h)
    old_c = np.concatenate([finals["fw"][0], finals["bw"][0]])   # model.py:117-118
    old_h = np.concatenate([finals["fw"][1], finals["bw"][1]])
    c0 = np.maximum(old_c @ W[N_RED % "w_reduce_c"] + W[N_RED % "bias_reduce_c"], F32(0))   # model.py:119
    h0 = np.maximum(old_h @ W[N_RED % "w_reduce_h"] + W[N_RED % "bias_reduce_h"], F32(0))   # model.py:120
    return out, (c0.astype(F32), h0.astype(F32))


def encoder_features(W, enc_states: np.ndarray) -> np.ndarray:
    """attention_decoder.py:66-67: 1x1 conv2d == matmul over the last axis, no bias."""
    A = enc_states.shape[-1]
    return (enc_states.reshape(-1, A) @ W[N_WH].reshape(A, A)).reshape(enc_states.shape).astype(F32)


def attention(W, enc_feat, enc_states, mask, c, h, coverage, use_coverage: bool):
    """attention_decoder.py:79-129.  enc_feat/enc_states [R,T,A], mask [R,T], c,h [R,H],
    coverage [R,T] or None.  Returns (context [R,A], attn [R,T], new_coverage)."""
    dec_feat = np.concatenate([c, h], axis=1) @ W[N_ATT_M] + W[N_ATT_B]         # :93, concat order (c,h) :203
    v = W[N_V]
    if use_coverage and coverage is not None:                                    # :103
        cov_feat = coverage[:, :, None] * W[N_WC].reshape(1, 1, -1)              # :105
        e = (v * np.tanh(enc_feat + dec_feat[:, None, :] + cov_feat, dtype=F32)).sum(axis=2, dtype=F32)   # :108
    else:
        e = (v * np.tanh(enc_feat + dec_feat[:, None, :], dtype=F32)).sum(axis=2, dtype=F32)              # :117
    a = softmax(e.astype(F32))                                                   # :98
    a = a * mask                                                                 # :99
    a = (a / a.sum(axis=1, keepdims=True, dtype=F32)).astype(F32)                # :100-101
    if use_coverage and coverage is not None:
        new_cov = (coverage + a).astype(F32)                                     # :114
    elif use_coverage:
        new_cov = a.copy()                                                       # :122-123
    else:
        new_cov = None
    ctx = (a[:, :, None] * enc_states).sum(axis=1, dtype=F32)                    # :126-127
    return ctx.astype(F32), a, new_cov


def final_dist(W, hps, out, p_gen, attn, ext_ids, n_oov):
    """model.py:228-238 (vocab softmax) + model.py:146-183 (pointer mixture).  out [R,H]."""
    scores = out @ W[N_PROJ_W] + W[N_PROJ_V]                    # model.py:236 xw_plus_b
    vocab_dist = softmax(scores.astype(F32))                    # model.py:238
    if not hps["pointer_gen"]:
        return vocab_dist
    R, V = vocab_dist.shape
    fin = np.zeros((R, V + n_oov), dtype=F32)
    fin[:, :V] = p_gen * vocab_dist                             # model.py:158,162-164
    copy = ((F32(1.0) - p_gen) * attn).astype(F32)              # model.py:159
    proj = np.zeros_like(fin)
    for r in range(R):                                          # tf.scatter_nd: duplicates accumulate (model.py:176)
        np.add.at(proj[r], ext_ids[r], copy[r])
    return (fin + proj).astype(F32)                             # model.py:181


def decode_onestep(W, hps, enc_states, enc_feat, mask, ext_ids, n_oov, tokens, c, h, prev_cov):
    """One sess.run of model.decode_onestep (model.py:367-443) = attention_decoder with
    initial_state_attention=True (attention_decoder.py:131-180) + projection + top-k.
    All row-batched over R = beam rows.  enc_feat may be None => recomputed (literal mode)."""
    use_cov = bool(hps["coverage"])
    if enc_feat is None:
        enc_feat = encoder_features(W, enc_states)              # attention_decoder.py:67 every step
    # :138-140 first pass: recompute previous context; this is what updates coverage
    ctx, _, cov = attention(W, enc_feat, enc_states, mask, c, h, prev_cov if use_cov else None, use_cov)
    inp = W[N_EMB][np.asarray(tokens, dtype=np.int64)]          # model.py:214
    x = np.concatenate([inp, ctx], axis=1) @ W[N_X_M] + W[N_X_B]                 # :150
    c2, h2 = lstm_cell(x.astype(F32), c, h, W[N_DEC_K], W[N_DEC_B])              # :153
    ctx2, attn, _ = attention(W, enc_feat, enc_states, mask, c2, h2, cov, use_cov)   # :156-158 (no cov update)
    p_gen = sigmoid(np.concatenate([ctx2, c2, h2, x], axis=1).astype(F32) @ W[N_PG_M] + W[N_PG_B])   # :166-167
    out = np.concatenate([h2, ctx2], axis=1) @ W[N_OUT_M] + W[N_OUT_B]           # :173 (cell_output == h)
    fin = final_dist(W, hps, out.astype(F32), p_gen.astype(F32), attn, ext_ids, n_oov)
    k2 = 2 * int(hps["batch_size"])                                              # model.py:284
    R = fin.shape[0]
    ids = np.zeros((R, k2), dtype=np.int32)
    logp = np.zeros((R, k2), dtype=F32)
    for r in range(R):
        p, i = top_k(fin[r], k2)
        ids[r, :len(i)] = i
        with np.errstate(divide="ignore"):
            logp[r, :len(i)] = np.log(p, dtype=F32)                              # model.py:285
    return ids, logp, c2, h2, attn, p_gen[:, 0], cov, ctx2


@dataclasses.dataclass
class Hyp:
    """beam_search.py:26-79."""
    tokens: List[int]
    log_probs: List[float]
    c: np.ndarray
    h: np.ndarray
    coverage: Optional[np.ndarray]
    ctx: Optional[np.ndarray] = None       # dedup mode only: carried context
    p_gens: List[float] = dataclasses.field(default_factory=list)

    @property
    def log_prob(self):                     # :72-74: Python sum over [0.0, f32, f32, ...] -> float64
        s = 0.0
        for lp in self.log_probs:
            s = s + float(lp)
        return s

    @property
    def avg_log_prob(self):                 # :77-79; len(tokens) counts [START]
        return self.log_prob / len(self.tokens)


def sort_hyps(hyps: List[Hyp]) -> List[Hyp]:
    """beam_search.py:171-173: Python's sort is stable, descending via reverse=True keeps the
    original order among equal keys."""
    return sorted(hyps, key=lambda x: x.avg_log_prob, reverse=True)


def run_beam_search(W, hps, enc_input: Sequence[int], enc_ext: Sequence[int], n_oov: int,
                    mode: str = "literal", trace: Optional[list] = None) -> Hyp:
    """beam_search.py:82-168.  mode="literal": K replicated rows, encoder features recomputed every
    step, two attention passes per step -- the work the reference does.  mode="dedup": encoder
    features computed once and the first attention pass of step t>=1 replaced by the carried
    context/attention of step t-1 (identical inputs => identical values)."""
    K = int(hps["beam_size"])
    V = W[N_EMB].shape[0]
    T = len(enc_input)
    use_cov = bool(hps["coverage"])
    enc_states1, (c0, h0) = run_encoder(W, np.asarray(enc_input))                # :95 (row 0 kept, model.py:363)
    feat1 = encoder_features(W, enc_states1[None])[0] if mode == "dedup" else None
    hyps = [Hyp([START_ID], [0.0], c0, h0, np.zeros((T,), dtype=F32) if use_cov else None)
            for _ in range(K)]                                                   # :100-106
    results: List[Hyp] = []
    steps = 0
    ext = np.asarray(enc_ext, dtype=np.int64)
    while steps < int(hps["max_dec_steps"]) and len(results) < K:                # :110
        R = len(hyps)
        latest = [t if t < V else UNK_ID for t in (hh.tokens[-1] for hh in hyps)]   # :111-113
        c = np.stack([hh.c for hh in hyps]).astype(F32)
        h = np.stack([hh.h for hh in hyps]).astype(F32)
        cov = np.stack([hh.coverage for hh in hyps]).astype(F32) if use_cov else None
        enc_states = np.broadcast_to(enc_states1, (R,) + enc_states1.shape)
        mask = np.ones((R, T), dtype=F32)                                        # batcher.py:175-176
        ext_r = np.broadcast_to(ext, (R, T))
        if mode == "literal":
            ids, logp, c2, h2, attn, pg, ncov, _ = decode_onestep(
                W, hps, enc_states, None, mask, ext_r, n_oov, latest, c, h, cov)
        else:
            feat = np.broadcast_to(feat1, (R,) + feat1.shape)
            ids, logp, c2, h2, attn, pg, ncov, nctx = _decode_onestep_dedup(
                W, hps, enc_states, feat, mask, ext_r, n_oov, latest, c, h, cov, hyps, steps)
        all_hyps: List[Hyp] = []
        num_orig = 1 if steps == 0 else len(hyps)                                # :127-129
        for i in range(num_orig):
            for j in range(2 * K):                                               # :133-141
                hh = hyps[i]
                all_hyps.append(Hyp(hh.tokens + [int(ids[i, j])], hh.log_probs + [logp[i, j]],
                                    c2[i], h2[i], ncov[i] if use_cov else None,
                                    ctx=(nctx[i], attn[i]) if mode == "dedup" else None,
                                    p_gens=hh.p_gens + [float(pg[i])]))
        hyps = []
        ordered = sort_hyps(all_hyps)
        if trace is not None:
            trace.append([(tuple(x.tokens[-2:]), x.log_prob) for x in ordered])
        for hh in ordered:                                                       # :145-154
            if hh.tokens[-1] == STOP_ID:
                if steps >= int(hps["min_dec_steps"]):
                    results.append(hh)
            else:
                hyps.append(hh)
            if len(hyps) == K or len(results) == K:
                break
        steps += 1                                                               # :156
    if len(results) == 0:                                                        # :160-162
        results = hyps
    return sort_hyps(results)[0]                                                 # :165-168


def _decode_onestep_dedup(W, hps, enc_states, feat, mask, ext_r, n_oov, latest, c, h, cov, hyps, steps):
    """Same outputs as decode_onestep, but the first attention pass is taken from the carried
    (context, attention) of the parent's previous step when steps >= 1."""
    use_cov = bool(hps["coverage"])
    if steps == 0:
        return decode_onestep(W, hps, enc_states, feat, mask, ext_r, n_oov, latest, c, h, cov)
    ctx = np.stack([hh.ctx[0] for hh in hyps]).astype(F32)
    a_prev = np.stack([hh.ctx[1] for hh in hyps]).astype(F32)
    ncov = (cov + a_prev).astype(F32) if use_cov else None                       # attention_decoder.py:114
    inp = W[N_EMB][np.asarray(latest, dtype=np.int64)]
    x = np.concatenate([inp, ctx], axis=1) @ W[N_X_M] + W[N_X_B]
    c2, h2 = lstm_cell(x.astype(F32), c, h, W[N_DEC_K], W[N_DEC_B])
    ctx2, attn, _ = attention(W, feat, enc_states, mask, c2, h2, ncov, use_cov)
    p_gen = sigmoid(np.concatenate([ctx2, c2, h2, x], axis=1).astype(F32) @ W[N_PG_M] + W[N_PG_B])
    out = np.concatenate([h2, ctx2], axis=1) @ W[N_OUT_M] + W[N_OUT_B]
    fin = final_dist(W, hps, out.astype(F32), p_gen.astype(F32), attn, ext_r, n_oov)
    k2 = 2 * int(hps["batch_size"])
    R = fin.shape[0]
    ids = np.zeros((R, k2), dtype=np.int32)
    logp = np.zeros((R, k2), dtype=F32)
    for r in range(R):
        p, i = top_k(fin[r], k2)
        ids[r, :len(i)] = i
        with np.errstate(divide="ignore"):
            logp[r, :len(i)] = np.log(p, dtype=F32)
    return ids, logp, c2, h2, attn, p_gen[:, 0], ncov, ctx2


def decode_ids(W, hps, enc_input, enc_ext, n_oov, mode="literal", trace=None):
    """Tensor-level decode of one article: returns (tokens after [START] incl. a trailing [STOP]
    if the hypothesis finished, per-token float32 log-probs, avg_log_prob)."""
    if len(enc_input) == 0:          # zero-token article => empty summary, no model work
        return [], [], 0.0
    best = run_beam_search(W, hps, enc_input, enc_ext, n_oov, mode=mode, trace=trace)
    return [int(t) for t in best.tokens[1:]], [F32(x) for x in best.log_probs[1:]], best.avg_log_prob


def decode_row(W, vocab: Vocab, hps, uuid: bytes, article: bytes, reference: bytes, mode="literal"):
    """decode.py:95-129 + write_for_flink :159-185 + flink_writer.py:25-34.  Output row
    (uuid, article, summary, reference).  The reference column is passed through unchanged (the
    reference re-tokenises it with nltk, batcher.py:579 -- third-party, documented deviation)."""
    ex = make_example(article, vocab, int(hps["max_enc_steps"]))
    toks, _, _ = decode_ids(W, hps, ex.enc_input, ex.enc_input_extend_vocab, len(ex.article_oovs), mode=mode)
    words = cut_at_stop(outputids2words(toks, vocab, ex.article_oovs))           # decode.py:109-118
    return uuid, article, format_summary(words), reference
