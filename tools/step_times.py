"""In-process per-kernel-group timing (CUDA events inside the library, option profile=1), no profiler attached.
Usage: python tools/step_times.py [articles_per_batch] [extra flags] [option sets separated by '/', options by ',']
e.g.   python tools/step_times.py 64 "" "attention=fused/attention=split/vocab=simt" """
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
extra = sys.argv[2] if len(sys.argv) > 2 else ""
sets = [x.split(",") if x else [] for x in (sys.argv[3].split("/") if len(sys.argv) > 3 else ["", "vocab=simt"])]
beam = int(os.environ.get("PGEN_BEAM", "4"))
paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
flags = synth.hyperparams_string(paths, beam_size=beam, extra="--articles_per_batch=%d --slots=1 %s" % (B, extra))
arts = synth.make_article_ids(2 * B, seed=1234)
with native.Decoder(flags) as dec:
    b0 = native.pack_articles(arts[:B], dec.max_enc_steps)
    for opt in sets:
        for o in opt: dec.set_option(o)
        dec.set_option("profile=0")
        for _ in range(2):
            dec.stage_upload(*b0); dec.stage_run(); dec.stage_sync()
        graph_ms = dec.stats()["ms_decoder"]
        dec.set_option("profile=1")
        dec.stage_upload(*b0); dec.stage_run(); dec.stage_sync()
        st = dec.stats()
        n = max(1.0, st["prof_steps"])
        dec.set_option("profile=0")
        dec.set_option("measure=30")
        pm = dec.stats()
        print(opt, "encoder %.3f ms decoder %.3f ms (graphs: %.3f) | per step: attention %.1f us, vocab gemm %.1f us, select+merge %.1f us (steps %d) | replayed alone: attention %.1f us, vocab gemm %.1f us" % (
            st["ms_encoder"], st["ms_decoder"], graph_ms, 1e3 * st["ms_attention"] / n, 1e3 * st["ms_vocab"] / n, 1e3 * st["ms_select"] / n, n,
            1e3 * pm["ms_attention"] / max(1.0, pm["prof_steps"]), 1e3 * pm["ms_vocab"] / max(1.0, pm["prof_steps"])))
