"""Synthetic vocabulary, articles and weights.

There is no network, so neither the CNN/DailyMail vocab nor the pretrained checkpoint
(log/download_model.sh:14-28) can be fetched.  Everything below is seeded and deterministic.

Weight sets
  W0 "reference-scale": the reference's own initialisers -- LSTM kernels U(-0.02, 0.02)
     (run_summarization.py:72, model.py:206), embedding / output projection / reduce weights
     truncated-normal sigma=1e-4 (:73, model.py:207), attention_decoder get_variable() calls
     Glorot-uniform (TF default), biases 0.  The 50k-way distribution is then almost flat,
     so W0 is used for log-prob tolerance only.
  W1 "trained-scale": same shapes with O(1) logits, used for token identity, early stop
     and near-tie accounting.
"""
from __future__ import annotations

import os
from typing import Dict, List, Tuple

import numpy as np

P = "seq2seq/"
D = P + "decoder/attention_decoder/"


def variable_shapes(vsize: int, hidden: int = 256, emb: int = 128) -> Dict[str, Tuple[int, ...]]:
    """TF variable names -> shapes of the decode graph (model.py:204-231, attention_decoder.py:51-73,
    :150-173, :218-227)."""
    H, E, A = hidden, emb, 2 * hidden
    sh = {P + "embedding/embedding": (vsize, E)}
    for d in ("fw", "bw"):
        sh[P + "encoder/bidirectional_rnn/%s/lstm_cell/kernel" % d] = (E + H, 4 * H)
        sh[P + "encoder/bidirectional_rnn/%s/lstm_cell/bias" % d] = (4 * H,)
    sh[P + "reduce_final_st/w_reduce_c"] = (A, H)
    sh[P + "reduce_final_st/w_reduce_h"] = (A, H)
    sh[P + "reduce_final_st/bias_reduce_c"] = (H,)
    sh[P + "reduce_final_st/bias_reduce_h"] = (H,)
    sh[D + "W_h"] = (1, 1, A, A)
    sh[D + "v"] = (A,)
    sh[D + "coverage/w_c"] = (1, 1, 1, A)
    sh[D + "Attention/Linear/Matrix"] = (2 * H, A)
    sh[D + "Attention/Linear/Bias"] = (A,)
    sh[D + "Linear/Matrix"] = (E + A, E)
    sh[D + "Linear/Bias"] = (E,)
    sh[D + "lstm_cell/kernel"] = (E + H, 4 * H)
    sh[D + "lstm_cell/bias"] = (4 * H,)
    sh[D + "calculate_pgen/Linear/Matrix"] = (A + H + H + E, 1)
    sh[D + "calculate_pgen/Linear/Bias"] = (1,)
    sh[D + "AttnOutputProjection/Linear/Matrix"] = (H + A, H)
    sh[D + "AttnOutputProjection/Linear/Bias"] = (H,)
    sh[P + "output_projection/w"] = (H, vsize)
    sh[P + "output_projection/v"] = (vsize,)
    return sh


def _trunc_normal(rng, shape, std):
    x = rng.standard_normal(shape)
    bad = np.abs(x) > 2.0
    while bad.any():                       # tf.truncated_normal re-draws beyond 2 sigma
        x[bad] = rng.standard_normal(int(bad.sum()))
        bad = np.abs(x) > 2.0
    return (x * std).astype(np.float32)


def _glorot(rng, shape):
    if len(shape) == 1:
        fan_in = fan_out = shape[0]
    elif len(shape) == 2:
        fan_in, fan_out = shape
    else:                                   # conv kernel [h, w, in, out]
        rf = int(np.prod(shape[:-2]))
        fan_in, fan_out = shape[-2] * rf, shape[-1] * rf
    lim = np.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def make_weights(kind: str = "W1", vsize: int = 50000, hidden: int = 256, emb: int = 128,
                 seed: int = 20240611) -> Dict[str, np.ndarray]:
    rng = np.random.default_rng(seed)
    shapes = variable_shapes(vsize, hidden, emb)
    if kind == "W0":
        lstm_mag, tn_std, glorot_gain, stop_bias, bias_std = 0.02, 1e-4, 1.0, 0.0, 0.0
        emb_std = proj_std = red_std = tn_std
    elif kind == "W1":
        lstm_mag, emb_std, proj_std, red_std = 0.16, 0.5, 0.7, 0.05   # 0.32 makes the 256-unit LSTMs chaotic
        glorot_gain, stop_bias, bias_std = 1.5, 13.0, 0.05
    else:
        raise ValueError("unknown weight set %r" % kind)
    W = {}
    for name in sorted(shapes):
        sh = shapes[name]
        leaf = name.rsplit("/", 1)[-1]
        if leaf in ("kernel",):
            W[name] = rng.uniform(-lstm_mag, lstm_mag, size=sh).astype(np.float32)
        elif name.endswith("embedding/embedding"):
            W[name] = _trunc_normal(rng, sh, emb_std)
        elif name.endswith("output_projection/w"):
            W[name] = _trunc_normal(rng, sh, proj_std)
        elif name.endswith("output_projection/v"):
            W[name] = _trunc_normal(rng, sh, 0.35 if kind == "W1" else emb_std)
        elif "reduce_final_st" in name:
            W[name] = _trunc_normal(rng, sh, red_std)
        elif leaf in ("bias", "Bias"):
            W[name] = (rng.standard_normal(sh) * bias_std).astype(np.float32)
        else:
            gain = glorot_gain
            if kind == "W1":            # sharper attention so that softmax/coverage/copy are all exercised
                gain *= 10.0 if leaf == "v" else 2.5 if leaf == "W_h" or "Attention/Linear/Matrix" in name else 1.0
            W[name] = (_glorot(rng, sh) * gain).astype(np.float32)
    if stop_bias:                       # tuned so that a good fraction of articles finish in (min, max)_dec_steps
        W[P + "output_projection/v"][3] = np.float32(stop_bias)      # [STOP] = id 3 (data.py:54)
    return W


def write_vocab_file(path: str, vsize: int = 50000) -> None:
    """49 996 lines ``w<rank> <count>`` for vsize=50000 so that Vocab reaches exactly vsize
    (data.py:51-73).  A few real-looking tokens ("." and ",") are placed at ranks 4 and 5 so that
    sentence splitting (decode.py:167-175) is exercised."""
    with open(path, "w") as f:
        for r in range(4, vsize):
            w = "." if r == 4 else "," if r == 5 else "w%d" % r
            f.write("%s %d\n" % (w, max(1, 10_000_000 // (r - 3))))


def vocab_word(rank: int) -> str:
    return "." if rank == 4 else "," if rank == 5 else "w%d" % rank


def make_article_ids(n: int, vsize: int = 50000, seed: int = 1234, max_len: int = 400,
                     oov_rate: float = 0.02, fixed_oovs: int = 0, full_len_only: bool = False
                     ) -> List[Tuple[np.ndarray, np.ndarray, int]]:
    """Token-id form of the synthetic articles: list of (enc_ids [T] with OOV->0, ext_ids [T] with the
    j-th distinct OOV -> vsize+j, n_oov).  Zipf(s=1.07) over ranks 4..vsize-1; lengths 84 % = max_len,
    15 % uniform in [50, max_len-1], 1 % of 1-5 tokens (the "article 0." shape,
    TensorFlowTest.java:209).  fixed_oovs > 0 lays out exactly that many distinct OOVs (config 4)."""
    rng = np.random.default_rng(seed)
    ranks = np.arange(4, vsize, dtype=np.float64)
    p = (ranks - 3.0) ** -1.07
    cdf = np.cumsum(p / p.sum())
    out = []
    for _ in range(n):
        u = rng.random()
        if full_len_only or u < 0.84:
            T = max_len
        elif u < 0.99:
            T = int(rng.integers(min(50, max_len), max_len)) if max_len > 50 else max_len
        else:
            T = int(rng.integers(1, 6))
        ids = (np.searchsorted(cdf, rng.random(T)) + 4).astype(np.int32)
        ids = np.minimum(ids, vsize - 1)
        ext = ids.copy()
        if fixed_oovs > 0 and T >= fixed_oovs:
            pos = rng.choice(T, size=min(T, max(fixed_oovs, int(T * 0.6))), replace=False)
            which = np.concatenate([np.arange(fixed_oovs), rng.integers(0, fixed_oovs, size=len(pos) - fixed_oovs)])
            order = np.argsort(pos)          # first appearance order must be 0,1,2,...
            pos_sorted = pos[order]
            lab = which[order]
            remap, nxt = {}, 0
            for q, l in zip(pos_sorted, lab):
                if int(l) not in remap:
                    remap[int(l)] = nxt
                    nxt += 1
                ext[q] = vsize + remap[int(l)]
                ids[q] = 0
            n_oov = nxt
        else:
            is_oov = rng.random(T) < oov_rate
            pool = max(1, int(is_oov.sum() * 0.7))
            remap, nxt = {}, 0
            for q in np.nonzero(is_oov)[0]:
                l = int(rng.integers(0, pool))
                if l not in remap:
                    remap[l] = nxt
                    nxt += 1
                ext[q] = vsize + remap[l]
                ids[q] = 0
            n_oov = nxt
        out.append((ids, ext, n_oov))
    return out


def ids_to_article(ext_ids: np.ndarray, vsize: int, salt: int = 0) -> bytes:
    """Render extended ids as the lower-case, single-space-separated text the reference's
    data prep produces (make_datafiles.py:121-142)."""
    words = []
    for t in ext_ids.tolist():
        words.append(vocab_word(t) if t < vsize else "oov%d_%d" % (salt, t - vsize))
    return " ".join(words).encode()


def make_articles(n: int, vsize: int = 50000, seed: int = 1234, **kw) -> List[bytes]:
    return [ids_to_article(ext, vsize, salt=i) for i, (_, ext, _) in
            enumerate(make_article_ids(n, vsize, seed, **kw))]


def make_model_dir(root: str, kind: str = "W1", vsize: int = 50000, hidden: int = 256, emb: int = 128,
                   exp_name: str = "synthetic_model") -> Dict[str, str]:
    """Lay out ``<root>/log/<exp_name>/train/model.pgenw`` and ``<root>/vocab`` the way the
    reference lays out its log_root (run_summarization.py:300, util.py:34)."""
    from .weights_io import save_weights, DEFAULT_NAME
    train = os.path.join(root, "log", exp_name, "train")
    os.makedirs(train, exist_ok=True)
    wpath = os.path.join(train, DEFAULT_NAME)
    if not os.path.exists(wpath):
        save_weights(wpath, make_weights(kind, vsize, hidden, emb))
    vpath = os.path.join(root, "vocab")
    if not os.path.exists(vpath):
        write_vocab_file(vpath, vsize)
    return {"vocab_path": vpath, "log_root": os.path.join(root, "log"), "exp_name": exp_name,
            "weights": wpath}


def hyperparams_string(paths: Dict[str, str], beam_size: int = 4, max_enc_steps: int = 400,
                       max_dec_steps: int = 100, min_dec_steps: int = 35, coverage: int = 1,
                       vocab_size: int = 50000, extra: str = "") -> str:
    """The space-joined flag string TFModel puts under its hyper-params key (TFModel.java:47;
    same flag list as TensorFlowTest.java:40-53)."""
    s = ("run_summarization.py --mode=decode --vocab_path=%s --log_root=%s --exp_name=%s "
         "--batch_size=4 --beam_size=%d --max_enc_steps=%d --max_dec_steps=%d --min_dec_steps=%d "
         "--coverage=%d --vocab_size=%d --single_pass=1 --inference=1"
         % (paths["vocab_path"], paths["log_root"], paths["exp_name"], beam_size, max_enc_steps,
            max_dec_steps, min_dec_steps, coverage, vocab_size))
    return s + (" " + extra if extra else "")
