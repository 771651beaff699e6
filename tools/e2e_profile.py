"""Where the end-to-end (submit/poll) time goes: host text path, launch calls, waiting for the GPU."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth
paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
flags = synth.hyperparams_string(paths, extra="--articles_per_batch=64 --flush_ms=-1 " + (sys.argv[1] if len(sys.argv) > 1 else ""))
NB = 16
ids = synth.make_article_ids(64 * NB, seed=5)
texts = [synth.ids_to_article(e, 50000, salt=i) for i, (_, e, _) in enumerate(ids)]
with native.Decoder(flags) as dec:
    for i in range(128): dec.submit(b"w", texts[i], b"")
    dec.poll(flush=True)
    s0 = dec.stats()
    t_sub = t_poll = 0.0
    t0 = time.perf_counter()
    n = 0
    for i, a in enumerate(texts):
        t1 = time.perf_counter(); dec.submit(b"u%d" % i, a, b"r"); t2 = time.perf_counter()
        n += len(dec.poll()); t3 = time.perf_counter()
        t_sub += t2 - t1; t_poll += t3 - t2
    t1 = time.perf_counter(); n += len(dec.poll(flush=True)); t_flush = time.perf_counter() - t1
    total = time.perf_counter() - t0
    s1 = dec.stats()
    print("rows %d total %.1f ms -> %.0f articles/s | in submit %.1f ms, in poll %.1f ms, final flush %.1f ms | library host-text %.1f ms | per batch: submit %.2f ms"
          % (n, total * 1e3, n / total, t_sub * 1e3, t_poll * 1e3, t_flush * 1e3, s1["ms_host_text"] - s0["ms_host_text"], t_sub * 1e3 / NB))
