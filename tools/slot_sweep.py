"""Device-resident throughput (articles/s) for several numbers of batches in flight and kernel variants.
Usage: python tools/slot_sweep.py "1,2,4" "decoder=fused,attention=fused/decoder=split,attention=split" """
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth
slots_list = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "1,2,4").split(",")]
optsets = [x.split(",") if x else [] for x in (sys.argv[2] if len(sys.argv) > 2 else "").split("/")]
B = 64
paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
arts = synth.make_article_ids(8 * B, seed=1234)
for slots in slots_list:
    flags = synth.hyperparams_string(paths, extra="--articles_per_batch=%d --slots=%d --flush_ms=-1" % (B, slots))
    with native.Decoder(flags) as dec:
        for opts in optsets:
            for o in opts: dec.set_option(o)
            batches = [native.pack_articles(arts[i * B:(i + 1) * B], dec.max_enc_steps) for i in range(8)]
            for rep in range(3):
                for sl in range(slots):
                    dec.stage_slot(sl); dec.stage_upload(*batches[(rep + sl) % 8]); dec.stage_run()
            for sl in range(slots):
                dec.stage_slot(sl); dec.stage_sync()
            torch.cuda.synchronize()
            n = 4 * slots
            t0 = time.perf_counter()
            for i in range(n):
                dec.stage_slot(i % slots); dec.stage_run()
            for sl in range(slots):
                dec.stage_slot(sl); dec.stage_sync()
            dt = time.perf_counter() - t0
            print("slots %d %-40s %7.0f articles/s  (%.2f ms per batch)" % (slots, ",".join(opts), n * B / dt, 1e3 * dt / n), flush=True)
            dec.stage_slot(0)
