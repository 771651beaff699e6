#!/usr/bin/env python
"""Headline benchmark: articles/s of beam-4 pointer-generator decode (400 encoder / 100 decoder steps).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]

One "step" = one pass of the hot path (encoder + beam search) over one device batch of 64 synthetic
CNN/DailyMail-shaped articles (BASELINE.json configs[2]).  Prints ONE JSON line on rank 0.

  value     articles/s with the batch's ids already in HBM when each timed interval starts
            (CUDA events on the library's stream around pgen_stage_run, summed over the K steps)
  e2e       the same metric through the operator-facing C ABI (pgen_submit / pgen_poll with article
            byte strings in host memory): tokenisation, H2D, decode, D2H and detokenisation inside
            the timed region
  roofline  the dominant kernel (pair) of the step, timed live: after a profiled extra batch the library replays the
            attention pair and the tensor-core pass 50 times each from a graph between two CUDA events on its
            stream (option measure=50); the per-step interval of the profiled batch is reported beside it
  cpu_baseline  the oracle in reference-faithful ("literal") mode on the host cores, bounded sample

--impl reference times that CPU restatement alone (TensorFlow 1.x, which the reference needs, is not
installable here: see DESIGN.md), on rank 0 only.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

import pgen_pkg  # noqa: E402

pgen_pkg.load()
from textsummarization_on_flink_b200 import synth  # noqa: E402

METRIC = "articles/sec beam-4 decode (400/100 steps)"
ARTICLES_PER_STEP = 64
BEAM, T_ENC, S_DEC, V, H, E = 4, 400, 100, 50000, 256, 128
MODEL_ROOT = os.environ.get("PGEN_BENCH_MODEL", os.path.join(tempfile.gettempdir(), "pgen_bench_model"))


def config_dict(n_gpus: int) -> dict:
    return {"workload": "BASELINE.json configs[2]: full beam-4 decode, 64 articles per device batch, "
                        "max_enc_steps=400, max_dec_steps=100, min_dec_steps=35, coverage=1, pointer_gen (OOV copy "
                        "scatter), vocab 50000, hidden 256, emb 128, W1 synthetic weights",
            "articles_per_step_per_gpu": ARTICLES_PER_STEP, "slots_in_flight": "library default (--slots=4) unless PGEN_BENCH_EXTRA overrides", "beam_size": BEAM, "max_enc_steps": T_ENC,
            "max_dec_steps": S_DEC, "vocab_size": V,
            "sharding": "articles sharded by rank, one GPU per rank, no collective on the data path (x%d)" % n_gpus,
            "l2": "no explicit flush: every decode step of every batch in flight streams its own 105 MB of encoder "
                  "states/features plus the 51 MB projection (and 32 MB of exact rows), and a batch's encoder reads "
                  "210 MB of the per-word table -- four batches in flight cycle > 600 MB through the 126 MB L2"}


def peaks() -> dict:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        d["source"] = "measured"
        return d
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu: int):
        self.gpu, self.proc, self.path = gpu, None, None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=f,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def ensure_model(rank: int, barrier) -> dict:
    if rank == 0:
        paths = synth.make_model_dir(MODEL_ROOT, "W1", vsize=V, hidden=H, emb=E)
    barrier()
    if rank != 0:
        paths = synth.make_model_dir(MODEL_ROOT, "W1", vsize=V, hidden=H, emb=E)
    return paths


def use_all_host_threads() -> int:
    """torchrun exports OMP_NUM_THREADS=1; the CPU arm is meant to use every host core."""
    n = os.cpu_count() or 1
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=n)
    except Exception:
        pass
    return n


def cpu_reference_sample(paths: dict, n_articles: int, seed: int) -> tuple:
    """The oracle in reference-faithful mode (K replicated rows, W_h.h_i recomputed every step, two
    attention passes, host-side beam search).  bench.py is one of the few places allowed to import it."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pgen_oracle as O
    from textsummarization_on_flink_b200 import weights_io
    W = weights_io.load_weights(paths["weights"])
    hps = O.parse_hyperparams(synth.hyperparams_string(paths))
    arts = synth.make_article_ids(n_articles, vsize=V, seed=seed, full_len_only=False)
    use_all_host_threads()
    t0 = time.perf_counter()
    for ids, ext, n_oov in arts:
        O.decode_ids(W, hps, ids, ext, n_oov, mode="literal")
    return time.perf_counter() - t0, n_articles


def run_reference(args, rank: int, world: int):
    if rank != 0:
        return
    paths = ensure_model(0, lambda: None)
    per_step = 2
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pgen_oracle as O
    from textsummarization_on_flink_b200 import weights_io
    W = weights_io.load_weights(paths["weights"])
    hps = O.parse_hyperparams(synth.hyperparams_string(paths))
    arts = synth.make_article_ids(per_step * (args.steps + args.warmup), vsize=V, seed=4321)
    use_all_host_threads()
    times = []
    for s in range(args.steps + args.warmup):
        t0 = time.perf_counter()
        for ids, ext, n_oov in arts[s * per_step:(s + 1) * per_step]:
            O.decode_ids(W, hps, ids, ext, n_oov, mode="literal")
        if s >= args.warmup:
            times.append(time.perf_counter() - t0)
    total = sum(times)
    value = per_step * args.steps / total
    cores = os.cpu_count()
    sample = ("%d articles per step of the same synthetic workload, one at a time, reference-faithful mode "
              "(NumPy fp32 restatement; TensorFlow 1.x unavailable)" % per_step)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "articles/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args.gpus),
        "cpu_baseline": {"value": value, "unit": "articles/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "articles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--cpu-articles", type=int, default=8, help="size of the bounded CPU-baseline sample")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from textsummarization_on_flink_b200 import native
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the native arm has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    paths = ensure_model(rank, barrier)
    flags = synth.hyperparams_string(paths, beam_size=BEAM, max_enc_steps=T_ENC, max_dec_steps=S_DEC, vocab_size=V,
                                     extra="--articles_per_batch=%d --flush_ms=-1 %s"
                                           % (ARTICLES_PER_STEP, os.environ.get("PGEN_BENCH_EXTRA", "")))
    dec = native.Decoder(flags, device=local_rank)
    for opt in filter(None, os.environ.get("PGEN_BENCH_OPTS", "").split(",")):   # experiments only, e.g. encoder=steps
        dec.set_option(opt)
    n_total = args.warmup + args.steps
    # every rank decodes its own shard of the synthetic stream (weak scaling: per-GPU work is fixed)
    art_ids = synth.make_article_ids(ARTICLES_PER_STEP * n_total, vsize=V, seed=1234 + 1000 * rank)
    batches = [native.pack_articles(art_ids[i * ARTICLES_PER_STEP:(i + 1) * ARTICLES_PER_STEP], T_ENC)
               for i in range(n_total)]

    # ---------------- device-resident: value ----------------
    # The handle keeps `slots` device batches in flight on separate streams.  Each slot's batch of ids is uploaded
    # BEFORE the timed region and stays resident; the K timed steps alternate between the slots, every step
    # recomputing the whole path (encoder + beam search) for its slot's 64 articles.
    n_slots = max(1, dec.flag("slots"))
    streams = []
    for sl in range(n_slots):
        dec.stage_slot(sl)
        streams.append(torch.cuda.ExternalStream(dec.stream_ptr(), device=dev))
    for i in range(args.warmup):
        for sl in range(n_slots):
            dec.stage_slot(sl)
            dec.stage_upload(*batches[(i * n_slots + sl) % n_total])
            dec.stage_run()
    for sl in range(n_slots):
        dec.stage_slot(sl)
        dec.stage_sync()
        dec.stage_upload(*batches[(args.warmup + sl) % n_total])
        dec.stage_sync()
    st0 = dec.stats()
    sampler = ClockSampler(local_rank)
    barrier()
    torch.cuda.synchronize()
    sampler.start()
    e0 = torch.cuda.Event(enable_timing=True)
    e0.record(streams[0])
    for i in range(args.steps):
        dec.stage_slot(i % n_slots)
        dec.stage_run()                       # asynchronous; consecutive runs of one slot serialise on its stream
    ends = []
    for sl in range(n_slots):
        e1 = torch.cuda.Event(enable_timing=True)
        e1.record(streams[sl])
        ends.append(e1)
    for sl in range(n_slots):
        dec.stage_slot(sl)
        dec.stage_sync()
    torch.cuda.synchronize()
    barrier()
    clocks = sampler.stop()
    st1 = dec.stats()
    dec.stage_slot(0)
    dev_ms = max(e0.elapsed_time(e1) for e1 in ends)
    dev_ms_max = max_over_ranks(dev_ms)
    n_articles_all = ARTICLES_PER_STEP * args.steps * world
    value = n_articles_all / (dev_ms_max / 1e3)
    launches = int(st1["kernel_launches"] - st0["kernel_launches"])

    # ---------------- end to end through submit/poll: e2e ----------------
    texts = [synth.ids_to_article(ext, V, salt=i) for i, (_, ext, _) in enumerate(art_ids)]
    for i in range(min(args.warmup, 2) * ARTICLES_PER_STEP):
        dec.submit(b"w%d" % i, texts[i], b"")
    dec.poll(flush=True)
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n_rows = 0
    for i in range(args.warmup * ARTICLES_PER_STEP, n_total * ARTICLES_PER_STEP):
        dec.submit(b"uuid-%d" % i, texts[i], b"reference %d." % i)
        n_rows += len(dec.poll())
    n_rows += len(dec.poll(flush=True))
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    assert n_rows == ARTICLES_PER_STEP * args.steps, n_rows
    e2e_value = n_articles_all / e2e_s
    h2d = ARTICLES_PER_STEP * (5 * T_ENC + 4) * 4
    d2h = ARTICLES_PER_STEP * (2 * S_DEC * 4 + 4 * 3 + 8)

    # ---------------- roofline of the dominant kernel (profiled extra step, CUDA events in the library) ----------
    dec.set_option("profile=1")
    dec.stage_upload(*batches[-1])
    dec.stage_run()
    dec.stage_sync()
    ps = dec.stats()
    steps_done = dec.debug_fetch("steps_done", ARTICLES_PER_STEP, np.int32)
    dec.set_option("profile=0")
    # kernel-only durations: the attention pair and the tensor-core pass of that batch replayed 50 times each from a
    # graph between two CUDA events on the library's stream (no launch gaps; every article active)
    step_ms_vocab = ps["ms_vocab"] / max(1.0, ps["prof_steps"])
    step_ms_attn = ps["ms_attention"] / max(1.0, ps["prof_steps"])
    dec.set_option("measure=50")
    pm = dec.stats()
    pk = peaks()
    n_prof = max(1.0, pm["prof_steps"])
    ms_vocab, ms_attn = pm["ms_vocab"] / n_prof, pm["ms_attention"] / n_prof
    R = ARTICLES_PER_STEP * BEAM
    lens = batches[-1][2].astype(np.float64)
    A = 2 * H
    attn_bytes_total = float(np.sum(2 * lens * A * 4 + 3 * BEAM * lens * 4)) * n_prof
    roof_attn = {"bound": "hbm", "kernel": "attn_energy + attn_context (per step)",
                 "achieved": attn_bytes_total / (pm["ms_attention"] * 1e-3) / 1e9 if pm["ms_attention"] > 0 else None,
                 "peak": pk["hbm_gbs"], "unit": "GB/s", "ms_per_launch": ms_attn,
                 "ms_per_launch_in_profiled_step": step_ms_attn,
                 # dram__bytes_read.sum + dram__bytes_write.sum of one attn_energy + attn_context launch pair from the
                 # ncu --set full capture summarised in profiles/r1b_summary.md (batch with mean length 372:
                 # 96.9 MB algorithmic): bytes per launch pair
                 "traffic": 96.60e6, "algorithmic_bytes_per_launch": attn_bytes_total / max(1.0, n_prof)}
    if roof_attn["achieved"]:
        roof_attn["frac"] = roof_attn["achieved"] / pk["hbm_gbs"]
    flops = 2.0 * R * H * V
    roof_vocab = {"bound": "tensor", "kernel": "vocab projection GEMM (per step)",
                  "achieved": flops / (ms_vocab * 1e-3) / 1e12 if ms_vocab > 0 else None,
                  "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s", "ms_per_launch": ms_vocab,
                  "ms_per_launch_in_profiled_step": step_ms_vocab,
                  "traffic": 51.91e6, "note": "tf32 operands (kind::tf32); peak is the measured bf16 figure, the tf32 "
                                               "tensor peak is half of it; traffic = dram bytes of one vocab_tc_kernel launch"}
    if roof_vocab["achieved"]:
        roof_vocab["frac"] = roof_vocab["achieved"] / pk["bf16_tflops_sustained"]
    dominant = roof_vocab if ms_vocab >= ms_attn else roof_attn
    dominant = dict(dominant, peak_source=pk.get("source"))

    out = {
        "metric": METRIC, "value": value, "unit": "articles/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(world),
        "e2e": {"value": e2e_value, "unit": "articles/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": "pgen_submit/pgen_poll with article byte strings"},
        "gpu_launches": launches, "clocks": clocks, "roofline": dominant,
        "roofline_other": roof_attn if dominant.get("kernel") == roof_vocab["kernel"] else roof_vocab,
        "stage_ms_last_batch": {"h2d": ps["ms_h2d"], "encoder": ps["ms_encoder"], "decoder": ps["ms_decoder"],
                                "d2h": ps["ms_d2h"], "decode_steps": ps["prof_steps"]},
    }
    if rank == 0:
        try:
            secs, n = cpu_reference_sample(paths, args.cpu_articles, seed=99)
            out["cpu_baseline"] = {"value": n / secs, "unit": "articles/s", "cores": os.cpu_count(), "kind": "port",
                                   "sample": "%d articles of the same synthetic workload decoded one at a time by the "
                                             "NumPy restatement in reference-faithful mode (TF1 unavailable), %.1f s"
                                             % (n, secs)}
        except Exception as ex:      # the CPU baseline must never take the GPU number down with it
            out["cpu_baseline"] = {"value": None, "unit": "articles/s", "cores": os.cpu_count(), "kind": "port",
                                   "sample": "failed: %r" % (ex,)}
        print(json.dumps(out))
    dec.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
