"""In-process per-kernel-group timing (CUDA events inside the library, option profile=1), no profiler attached."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
extra = sys.argv[2] if len(sys.argv) > 2 else ""
paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
flags = synth.hyperparams_string(paths, extra="--articles_per_batch=%d --slots=1 %s" % (B, extra))
arts = synth.make_article_ids(2 * B, seed=1234)
with native.Decoder(flags) as dec:
    b0 = native.pack_articles(arts[:B], dec.max_enc_steps)
    for opt in ([], ["vocab=simt"]):
        for o in opt: dec.set_option(o)
        dec.set_option("profile=0")
        for _ in range(2):
            dec.stage_upload(*b0); dec.stage_run(); dec.stage_sync()
        dec.set_option("profile=1")
        dec.stage_upload(*b0); dec.stage_run(); dec.stage_sync()
        st = dec.stats()
        n = max(1.0, st["prof_steps"])
        print(opt, "encoder %.3f ms decoder %.3f ms | per step: attention %.1f us, vocab gemm %.1f us, select+merge %.1f us (steps %d)" % (
            st["ms_encoder"], st["ms_decoder"], 1e3 * st["ms_attention"] / n, 1e3 * st["ms_vocab"] / n, 1e3 * st["ms_select"] / n, n))
