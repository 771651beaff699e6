"""BASELINE.json configs[3] (beam 8, max_dec_steps 120, 200 OOVs per article) on one GPU: device-resident throughput
with the library's default number of batches in flight.  A parity-test configuration (tests/test_gpu_parity.py::
test_config4_beam8_200_oovs), timed here for the record; bench.py stays on configs[2]."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth

B, STEPS, WARM = 64, 12, 4
paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
flags = synth.hyperparams_string(paths, beam_size=8, max_dec_steps=120, extra="--articles_per_batch=%d --flush_ms=-1" % B)
arts = synth.make_article_ids(B * (STEPS + WARM), seed=77, fixed_oovs=200)
with native.Decoder(flags) as dec:
    n_slots = max(1, dec.flag("slots"))
    batches = [native.pack_articles(arts[i * B:(i + 1) * B], dec.max_enc_steps) for i in range(STEPS + WARM)]
    for i in range(WARM):
        dec.stage_slot(i % n_slots); dec.stage_upload(*batches[i]); dec.stage_run()
    for sl in range(n_slots):
        dec.stage_slot(sl); dec.stage_sync(); dec.stage_upload(*batches[WARM + sl]); dec.stage_sync()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(STEPS):
        dec.stage_slot(i % n_slots); dec.stage_run()
    for sl in range(n_slots):
        dec.stage_slot(sl); dec.stage_sync()
    dt = time.perf_counter() - t0
    st = dec.stats()
    print("config 4: %d articles in %.1f ms -> %.0f articles/s (%d slots, %.1f ms per batch of %d)" % (
        B * STEPS, dt * 1e3, B * STEPS / dt, n_slots, dt * 1e3 / STEPS, B))
