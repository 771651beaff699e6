import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth
B = 64
paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
flags = synth.hyperparams_string(paths, max_dec_steps=12, min_dec_steps=11, extra="--articles_per_batch=%d --slots=1" % B)
arts = synth.make_article_ids(B, seed=1234)
with native.Decoder(flags) as dec:
    b0 = native.pack_articles(arts, dec.max_enc_steps)
    for _ in range(2):
        dec.stage_upload(*b0); dec.stage_run(); dec.stage_sync()
    out = np.zeros(512 * 12, np.int64)
    rc = native.lib().pgen_debug_vfdbg(out.ctypes.data_as(C.c_void_p))
    d = out.reshape(512, 12)[:256]
    names = ["0 loads", "1a supergroups", "1b groups", "2 rescore", "3+4a dist", "4b merge+store", "4b rank count", "4b sync"]
    for i, n in enumerate(names):
        print("  %-16s mean %8.0f  max %8.0f  min %8.0f" % (n, d[:, i].mean(), d[:, i].max(), d[:, i].min()))
    print("  total mean %.0f max %.0f" % (d[:, :8].sum(1).mean(), d[:, :8].sum(1).max()))
