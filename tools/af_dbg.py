import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth
B = 64
paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
flags = synth.hyperparams_string(paths, max_dec_steps=12, min_dec_steps=5, extra="--articles_per_batch=%d --slots=1" % B)
arts = synth.make_article_ids(B, seed=1234)
with native.Decoder(flags) as dec:
    b0 = native.pack_articles(arts, dec.max_enc_steps)
    for _ in range(2):
        dec.stage_upload(*b0); dec.stage_run(); dec.stage_sync()
    dec.set_option("measure=5")        # all articles active, attention replayed alone
    print("measured attention us:", 1e3 * dec.stats()["ms_attention"] / 5)
    out = np.zeros(160 * 24, np.int64)
    rc = native.lib().pgen_debug_afdbg(out.ctypes.data_as(C.c_void_p))
    d = out.reshape(160, 24)[:148]
    en = ["loop", "get item", "wait ds", "wait F", "E phase", "wait efree", "sync+arrive", "-", "-", "-", "items", "total"]
    cn = ["loop", "get item", "wait efull", "softmax", "wait H", "C phase", "merge", "drain", "-", "merges", "items", "total"]
    for nm, off in ((en, 0), (cn, 12)):
        print("E warps (warp 0 lane 0):" if off == 0 else "C warps (first C warp lane 0):")
        for i, n in enumerate(nm):
            if n != "-":
                print("  %-12s mean %8.0f  max %8.0f  min %8.0f" % (n, d[:, off + i].mean(), d[:, off + i].max(), d[:, off + i].min()))
