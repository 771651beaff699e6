"""Debug helper (GPU box): step-by-step comparison of the device beam trace with the oracle's."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth, weights_io
import pgen_oracle as O

paths = synth.make_model_dir("/tmp/pgen_dbg_model", "W1")
flags = synth.hyperparams_string(paths, extra="--articles_per_batch=8")
hps = O.parse_hyperparams(flags)
W = weights_io.load_weights(paths["weights"])
arts = synth.make_article_ids(6, seed=77)
K, S = 4, 100
with native.Decoder(flags) as dec:
    ids, ext, lens, noov = native.pack_articles(arts, dec.max_enc_steps)
    tok, ln, lp, avg = dec.decode_ids(ids, ext, lens, noov)
    n = len(arts)
    ht = dec.debug_fetch("trace_tokens", n * S * K, np.int32).reshape(n, S, K)
    hp = dec.debug_fetch("trace_parents", n * S * K, np.int32).reshape(n, S, K)
    hl = dec.debug_fetch("trace_logp", n * S * K, np.float32).reshape(n, S, K)
for b, (a, e, k) in enumerate(arts):
    tr = []
    rt, rl, ravg = O.decode_ids(W, hps, a, e, k, mode="dedup", trace=tr)
    got = tok[b, :ln[b]].tolist()
    print("article", b, "len", len(a), "match", got == rt, "steps oracle", len(tr))
    # reconstruct oracle beams per step
    sums_gpu = np.zeros(K)
    for s, cand in enumerate(tr):
        sel = [c for c in cand if c[0][1] != 3][:K]
        otoks = [c[0][1] for c in sel]; osum = [c[1] for c in sel]
        gt = ht[b, s].tolist()
        par = hp[b, s]
        sums_gpu = np.array([sums_gpu[par[j]] + float(hl[b, s, j]) for j in range(K)])
        if gt != otoks:
            print("  first beam divergence at step", s, "gpu", gt, "oracle", otoks)
            print("  gpu sums", sums_gpu, " oracle sums", osum)
            print("  oracle sorted cands (top 12):", [(c[0], round(c[1], 6)) for c in cand[:12]])
            break
        d = np.abs(sums_gpu - np.array(osum))
        if s % 10 == 0 or d.max() > 1e-3:
            print("  step", s, "max |sum diff|", d.max())
