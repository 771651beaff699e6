"""Device-resident throughput of the configs[2] decode for other device-batch sizes / batches in flight (bench.py stays
on 64 articles per batch).  Usage: python tools/batch_sweep.py "128:2,128:4,32:8" """
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth

paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
for spec in (sys.argv[1] if len(sys.argv) > 1 else "64:4").split(","):
    B, slots = (int(x) for x in spec.split(":"))
    steps, warm = max(8, 1024 // B), slots
    flags = synth.hyperparams_string(paths, extra="--articles_per_batch=%d --slots=%d --flush_ms=-1" % (B, slots))
    arts = synth.make_article_ids(B * (steps + warm + slots), seed=1234)
    with native.Decoder(flags) as dec:
        batches = [native.pack_articles(arts[i * B:(i + 1) * B], dec.max_enc_steps) for i in range(steps + warm + slots)]
        for i in range(warm):
            dec.stage_slot(i % slots); dec.stage_upload(*batches[i]); dec.stage_run()
        for sl in range(slots):
            dec.stage_slot(sl); dec.stage_sync(); dec.stage_upload(*batches[warm + sl]); dec.stage_sync()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(steps):
            dec.stage_slot(i % slots); dec.stage_run()
        for sl in range(slots):
            dec.stage_slot(sl); dec.stage_sync()
        dt = time.perf_counter() - t0
    print("batch %4d x %d slots: %.0f articles/s (%.2f ms per batch)" % (B, slots, B * steps / dt, dt * 1e3 / steps), flush=True)
