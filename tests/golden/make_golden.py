"""Generates tests/golden/decode_small.npz and decode_full.npz from the ORACLE (the reference holds no golden
vectors and TensorFlow 1.x cannot run here: SURVEY.md section 8c).  They are not TensorFlow outputs; the same inputs run
through the reference's own Python sources (make_reference_golden.py -> ref_small.npz / ref_full.npz) give identical
token ids and log-probs within 1e-5 / 5e-4 (tests/test_reference_golden.py).   Run:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pgen_pkg  # noqa: E402

pgen_pkg.load()
import pgen_oracle as O  # noqa: E402
from textsummarization_on_flink_b200 import synth  # noqa: E402

SMALL = dict(V=300, H=32, E=16, seed=7, stop_bias=4.0,
             flags="p --mode=decode --beam_size=4 --coverage=1 --max_enc_steps=24 --max_dec_steps=30 --min_dec_steps=8",
             art_seed=100, n=12, max_len=24, oov_rate=0.15)
FULL = dict(flags="p --mode=decode --beam_size=4 --coverage=1 --max_enc_steps=400 --max_dec_steps=100 --min_dec_steps=35",
            art_seed=77, n=4)


def small_weights():
    W = synth.make_weights("W1", vsize=SMALL["V"], hidden=SMALL["H"], emb=SMALL["E"], seed=SMALL["seed"])
    W["seq2seq/output_projection/v"][3] = np.float32(SMALL["stop_bias"])
    return W


def decode_all(W, flags, arts):
    hps = O.parse_hyperparams(flags)
    S = int(hps["max_dec_steps"])
    tok = np.zeros((len(arts), S), np.int32)
    lp = np.zeros((len(arts), S), np.float32)
    ln = np.zeros((len(arts),), np.int32)
    avg = np.zeros((len(arts),), np.float64)
    for b, (a, e, k) in enumerate(arts):
        t, l, av = O.decode_ids(W, hps, a, e, k, mode="literal")
        ln[b] = len(t)
        tok[b, :len(t)] = t
        lp[b, :len(t)] = l
        avg[b] = av
    return dict(tokens=tok, lens=ln, logprobs=lp, avg=avg)


if __name__ == "__main__":
    arts = synth.make_article_ids(SMALL["n"], vsize=SMALL["V"], seed=SMALL["art_seed"], max_len=SMALL["max_len"], oov_rate=SMALL["oov_rate"])
    np.savez_compressed(os.path.join(HERE, "decode_small.npz"), **decode_all(small_weights(), SMALL["flags"], arts))
    arts = synth.make_article_ids(FULL["n"], seed=FULL["art_seed"])
    np.savez_compressed(os.path.join(HERE, "decode_full.npz"), **decode_all(synth.make_weights("W1"), FULL["flags"], arts))
    print("wrote", os.listdir(HERE))
