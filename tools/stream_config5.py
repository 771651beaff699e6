"""BASELINE.json configs[4]: "Flink ML pipeline streaming TFModel.transform over 100k synthetic articles, parallelism 8
(one GPU per subtask)" -- an EMULATION of TFModel.transform at that parallelism, no JVM and no Flink here.

What a TaskManager does with the operator (src/main/java/.../TFModel.java:55-81, param/HasClusterConfig.java:20-24,
src/test/java/.../TensorFlowTest.java:93-107,204-214) is replayed by P feeder THREADS of one process: each opens its own
handle on GPU (subtask index mod visible GPUs), then per row calls submit + poll on its own thread, flushes at end of
input and closes -- open -> N x flatMap -> close.  Rows are dealt round-robin to the subtasks like a rebalanced stream.

    python tools/stream_config5.py [--rows 100000] [--parallelism 8] [--flush-ms 50] [--articles-per-batch 64]

Prints one JSON line: rows/s over all subtasks (wall clock from the first submit to the last row out), per-row latency
(submit -> the poll that returned it) p50 / p99, mean batch fill, rows per GPU.
"""
import argparse
import faulthandler
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

import pgen_pkg  # noqa: E402

pgen_pkg.load()
from textsummarization_on_flink_b200 import native, sharding, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=100000)
    ap.add_argument("--parallelism", type=int, default=8)
    ap.add_argument("--flush-ms", type=int, default=50)
    ap.add_argument("--articles-per-batch", type=int, default=64)
    ap.add_argument("--pool", type=int, default=2048, help="distinct synthetic articles (cycled)")
    ap.add_argument("--options", default="", help="comma-separated pgen_set_option strings applied to every handle")
    ap.add_argument("--jitter-ms", type=float, default=0.0, help="a feeder sleeps up to this long before 1 row in 100 (a bursty source)")
    ap.add_argument("--spin-threads", type=int, default=0, help="extra Python threads that only compete for the interpreter lock")
    ap.add_argument("--watchdog-s", type=int, default=0, help="dump every thread's Python stack to stderr after this many seconds")
    args = ap.parse_args()
    handles = {}
    if args.watchdog_s > 0:
        faulthandler.dump_traceback_later(args.watchdog_s, repeat=True)

        def watchdog():                                                 # a run that does not end: say where every handle is
            import subprocess
            time.sleep(args.watchdog_s + 1)
            for rep in range(2):
                for k, dec in sorted(handles.items()):
                    print("[watchdog] subtask %d (GPU %d): %s" % (k, k % n_gpu, dec.debug_where()), file=sys.stderr, flush=True)
                print(subprocess.run(["nvidia-smi", "--query-gpu=index,utilization.gpu,clocks.sm,power.draw", "--format=csv,noheader"],
                                     capture_output=True, text=True).stdout, file=sys.stderr, flush=True)
                time.sleep(5)

        threading.Thread(target=watchdog, daemon=True).start()
    t_prog = time.perf_counter()

    def progress(msg):
        print("[%7.2f s] %s" % (time.perf_counter() - t_prog, msg), file=sys.stderr, flush=True)
    n_gpu = native.device_count()
    if n_gpu < 1:
        raise SystemExit("stream_config5: no CUDA device (the library has no CPU fallback)")
    paths = synth.make_model_dir(os.environ.get("PGEN_BENCH_MODEL", "/tmp/pgen_bench_model"), "W1")
    progress("model files ready, %d GPUs" % n_gpu)
    flags = synth.hyperparams_string(paths, extra="--articles_per_batch=%d --flush_ms=%d" % (args.articles_per_batch, args.flush_ms))
    pool = synth.make_articles(args.pool, seed=555)
    P = args.parallelism
    lat = [[] for _ in range(P)]
    stats = [None] * P
    errors = []
    ready = threading.Barrier(P + 1)
    t_first = [None] * P
    t_last = [None] * P

    def subtask(k):
        try:
            rows = sharding.shard_rows(range(args.rows), k, P)          # rebalance: row i -> subtask i mod P
            with native.Decoder(flags, device=k % n_gpu) as dec:        # open()
                handles[k] = dec
                for o in [x for x in args.options.split(",") if x]:
                    dec.set_option(o)
                rng = np.random.default_rng(1000 + k)
                progress("subtask %d: handle open on GPU %d" % (k, k % n_gpu))
                for i in range(2 * args.articles_per_batch):            # warm-up: plain launches, then the graph capture
                    dec.submit(b"w", pool[i % len(pool)], b"")
                dec.poll(flush=True)
                st0 = dec.stats()
                progress("subtask %d: warm" % k)
                ready.wait()
                t_in = {}
                t_first[k] = time.perf_counter()
                for i in rows:                                          # flatMap(row, out)
                    if args.jitter_ms > 0 and rng.random() < 0.01:
                        time.sleep(rng.random() * args.jitter_ms * 1e-3)
                    t_in[i] = time.perf_counter()
                    dec.submit(b"%d" % i, pool[i % len(pool)], b"")
                    out = dec.poll()
                    if out:
                        now = time.perf_counter()
                        lat[k].extend(now - t_in.pop(int(r[0])) for r in out)
                progress("subtask %d: end of input" % k)
                out = dec.poll(flush=True)                              # close(): end of input
                handles.pop(k, None)
                now = time.perf_counter()
                lat[k].extend(now - t_in.pop(int(r[0])) for r in out)
                t_last[k] = now
                st1 = dec.stats()
                assert not t_in, "%d rows never came back" % len(t_in)
                stats[k] = {"rows": st1["rows_out"] - st0["rows_out"], "batches": st1["batches"] - st0["batches"]}
        except Exception as ex:
            errors.append(repr(ex))
            try:
                ready.abort()
            except Exception:
                pass

    stop_spin = threading.Event()

    def spin():
        x = 0
        while not stop_spin.is_set():
            x += 1

    for _ in range(args.spin_threads):
        threading.Thread(target=spin, daemon=True).start()
    threads = [threading.Thread(target=subtask, args=(k,)) for k in range(P)]
    for t in threads:
        t.start()
    ready.wait()
    for t in threads:
        t.join()
    stop_spin.set()
    if errors:
        raise SystemExit("stream_config5: " + "; ".join(errors))
    wall = max(t_last) - min(t_first)
    all_lat = np.sort(np.concatenate([np.asarray(x) for x in lat]))
    rows = sum(s["rows"] for s in stats)
    batches = sum(s["batches"] for s in stats)
    per_gpu = [0] * n_gpu
    for k, s in enumerate(stats):
        per_gpu[k % n_gpu] += s["rows"]
    print(json.dumps({
        "what": "emulation of TFModel.transform at parallelism %d (feeder threads of one process, no JVM / Flink)" % P,
        "rows": rows, "parallelism": P, "gpus": n_gpu, "rows_per_s": rows / wall, "wall_s": wall,
        "latency_ms": {"p50": 1e3 * float(all_lat[len(all_lat) // 2]), "p99": 1e3 * float(all_lat[int(len(all_lat) * 0.99)]),
                       "max": 1e3 * float(all_lat[-1])},
        "mean_batch_fill": rows / max(1, batches), "articles_per_batch": args.articles_per_batch, "flush_ms": args.flush_ms,
        "rows_per_gpu": per_gpu}))


if __name__ == "__main__":
    main()
