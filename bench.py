#!/usr/bin/env python
"""Headline benchmark: articles/s of beam-4 pointer-generator decode (400 encoder / 100 decoder steps).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]

One "step" = one pass of the hot path (encoder + beam search) over one device batch of 64 synthetic
CNN/DailyMail-shaped articles (BASELINE.json configs[2]).  Prints ONE JSON line on rank 0.

  value     articles/s with the batch's ids already in HBM when each timed interval starts
            (CUDA events on the library's streams around pgen_stage_run, max over ranks)
  e2e       the same metric through the operator-facing C ABI (pgen_submit / pgen_poll with article
            byte strings in host memory): tokenisation, H2D, decode, D2H and detokenisation inside
            the timed region; `verified_rows` of those rows are then compared with the oracle
  roofline  the dominant kernel of the decode step (the attention pass), timed live: its interval in every step of
            one extra batch run with plain launches and CUDA events inside the library (option profile=1), against the
            measured HBM peak; `roofline_other` is the tensor-core vocabulary pass against a TF32 dense peak measured
            here with the same recipe MEASURED_PEAKS.json used for bf16
  config2 / config4   BASELINE.json configs[1] (encoder only) and configs[3] (beam 8, 120 steps, 200 OOVs) at this N
  cpu_baseline  the reference's algorithm restated for the CPU (PyTorch fp32, every host thread; TensorFlow 1.x,
            which the reference needs, cannot be installed here: DESIGN.md), bounded sample, rank 0 only

--impl reference times that CPU restatement alone, on rank 0's host cores whatever N is.
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
from textsummarization_on_flink_b200 import sharding, synth  # noqa: E402

METRIC = "articles/sec beam-4 decode (400/100 steps)"
ARTICLES_PER_STEP = 64
BEAM, T_ENC, S_DEC, V, H, E = 4, 400, 100, 50000, 256, 128
MODEL_ROOT = os.environ.get("PGEN_BENCH_MODEL", os.path.join(tempfile.gettempdir(), "pgen_bench_model"))


def config_dict(n_gpus: int) -> dict:
    return {"workload": "BASELINE.json configs[2]: full beam-4 decode, 64 articles per device batch, "
                        "max_enc_steps=400, max_dec_steps=100, min_dec_steps=35, coverage=1, pointer_gen (OOV copy "
                        "scatter), vocab 50000, hidden 256, emb 128, W1 synthetic weights",
            "articles_per_step_per_gpu": ARTICLES_PER_STEP, "slots_in_flight": "4 device batches in flight on 4 streams (library default --slots=4; --decoder=auto then runs the split-K decoder GEMMs + the one-launch attention); PGEN_BENCH_EXTRA overrides", "beam_size": BEAM, "max_enc_steps": T_ENC,
            "max_dec_steps": S_DEC, "vocab_size": V,
            "sharding": "articles sharded by rank, one GPU per rank, no collective on the data path (x%d)" % n_gpus,
            "l2": "no explicit flush: every decode step of every batch in flight streams its own 105 MB of encoder "
                  "states/features plus the 51 MB projection (and 32 MB of exact rows), and a batch's encoder reads "
                  "210 MB of the per-word table -- four batches in flight cycle > 600 MB through the 126 MB L2; "
                  "the roofline replays write a 256 MB buffer between replays (flush time subtracted)"}


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
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pgen_oracle_torch as OT
    n = OT.use_all_threads()
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=os.cpu_count() or 1)
    except Exception:
        pass
    return n


CPU_KIND_NOTE = ("PyTorch-CPU fp32 restatement of the reference algorithm (TensorFlow 1.x unavailable), one article at a "
                 "time, K replicated rows, W_h.h_i recomputed every step, two attention passes, host-side beam search")


def cpu_arm(paths: dict, arts, mode: str) -> float:
    """Seconds to decode `arts` with oracle/pgen_oracle_torch.py (bench.py is one of the few places allowed to import
    it).  mode "literal" = the work the reference does; "dedup" = without its redundancy (features once, one attention
    pass per step), so that the speed-up is not inflated by it."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import pgen_oracle as O
    import pgen_oracle_torch as OT
    from textsummarization_on_flink_b200 import weights_io
    W = weights_io.load_weights(paths["weights"])
    hps = O.parse_hyperparams(synth.hyperparams_string(paths))
    t0 = time.perf_counter()
    for ids, ext, n_oov in arts:
        OT.decode_ids(W, hps, ids, ext, n_oov, mode=mode, dtype=torch.float32)
    return time.perf_counter() - t0


def run_reference(args, rank: int, world: int):
    if rank != 0:
        return
    paths = ensure_model(0, lambda: None)
    per_step = 2
    threads = use_all_host_threads()
    arts = synth.make_article_ids(per_step * (args.steps + args.warmup) + 2, vsize=V, seed=4321)
    times = []
    for s in range(args.steps + args.warmup):
        secs = cpu_arm(paths, arts[s * per_step:(s + 1) * per_step], "literal")
        if s >= args.warmup:
            times.append(secs)
    total = sum(times)
    value = per_step * args.steps / total
    dedup = 2 / cpu_arm(paths, arts[-2:], "dedup")
    sample = "%d articles per step of the same synthetic workload; %s" % (per_step, CPU_KIND_NOTE)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "articles/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args.gpus),
        "note": "the CPU arm runs once, on rank 0's host cores, whatever --gpus says: it is not multiplied by N",
        "cpu_baseline": {"value": value, "unit": "articles/s", "cores": threads, "kind": "port", "sample": sample,
                         "dedup_value": dedup},
        "e2e": {"value": value, "unit": "articles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def measure_tf32_peak(torch, dev) -> dict:
    """Dense TF32 GEMM rate of this GPU the way MEASURED_PEAKS.json measured bf16: torch.matmul 8192^3 fp32 with TF32
    allowed (cuBLAS), best of 10 (burst) and back to back for ~1.5 s (sustained)."""
    n = 8192
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(n, n, device=dev)
        b = torch.randn(n, n, device=dev)
        for _ in range(3):
            a @ b
        torch.cuda.synchronize(dev)
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            e1.synchronize()
            best = min(best, e0.elapsed_time(e1))
        reps = max(10, int(1500.0 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            a @ b
        e1.record()
        e1.synchronize()
        sustained = e0.elapsed_time(e1) / reps
        fl = 2.0 * n ** 3
        return {"tf32_tflops": fl / best / 1e9, "tf32_tflops_sustained": fl / sustained / 1e9,
                "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS): best of 10 and %d back to back" % reps}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
        del a, b
        torch.cuda.empty_cache()


def profile_notes() -> dict:
    """Numbers that come from an ncu capture, not from this run: read from profiles/r2_ncu.json (which names the
    capture and its date), never typed into this file."""
    p = os.path.join(ROOT, "profiles", "r2_ncu.json")
    try:
        return json.load(open(p))
    except Exception:
        return {}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--cpu-articles", type=int, default=8, help="size of the bounded CPU-baseline sample")
    ap.add_argument("--verify", type=int, default=8, help="e2e rows compared with the oracle after the timed regions")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    if os.environ.get("PGEN_BENCH_WATCHDOG_S"):        # a run that does not end: every thread's Python stack on stderr
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["PGEN_BENCH_WATCHDOG_S"]), repeat=False)
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
    # whole job: every rank's articles over the slowest rank's device time (no collective on the data path)
    value = sharding.job_throughput(ARTICLES_PER_STEP * args.steps, dev_ms / 1e3, dist if world > 1 else None, dev)
    launches = int(st1["kernel_launches"] - st0["kernel_launches"])

    # ---------------- end to end through submit/poll: e2e ----------------
    texts = [synth.ids_to_article(ext, V, salt=i) for i, (_, ext, _) in enumerate(art_ids)]
    for i in range(min(args.warmup, 2) * ARTICLES_PER_STEP):
        dec.submit(b"w%d" % i, texts[i], b"")
    dec.poll(flush=True)
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e2e_rows = []
    for i in range(args.warmup * ARTICLES_PER_STEP, n_total * ARTICLES_PER_STEP):
        dec.submit(b"uuid-%d" % i, texts[i], b"reference %d." % i)
        e2e_rows += dec.poll()
    e2e_rows += dec.poll(flush=True)
    n_rows = len(e2e_rows)
    torch.cuda.synchronize()
    e2e_local_s = time.perf_counter() - t0
    e2e_s = max_over_ranks(e2e_local_s)
    barrier()
    assert n_rows == ARTICLES_PER_STEP * args.steps, n_rows
    e2e_value = sharding.job_throughput(ARTICLES_PER_STEP * args.steps, e2e_local_s, dist if world > 1 else None, dev)
    h2d = ARTICLES_PER_STEP * (5 * T_ENC + 4) * 4
    d2h = ARTICLES_PER_STEP * (2 * S_DEC * 4 + 4 * 3 + 8)

    # ---------------- verification: e2e rows of the timed region against the oracle ----------------
    verified = 0
    if rank == 0 and args.verify > 0:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import pgen_oracle as O
        from textsummarization_on_flink_b200 import weights_io
        Wv = weights_io.load_weights(paths["weights"])
        hpsv = O.parse_hyperparams(flags)
        vocab = O.Vocab(paths["vocab_path"], V)
        first = args.warmup * ARTICLES_PER_STEP
        near = 0
        for i in range(first, first + args.verify):
            uuid, article, summary, reference = e2e_rows[i - first]
            assert uuid == b"uuid-%d" % i and article == texts[i], "e2e row %d out of order" % i
            ex = O.make_example(texts[i], vocab, T_ENC)
            tr = []
            toks, _, _ = O.decode_ids(Wv, hpsv, ex.enc_input, ex.enc_input_extend_vocab, len(ex.article_oovs), mode="dedup", trace=tr)
            want = O.format_summary(O.cut_at_stop(O.outputids2words(toks, vocab, ex.article_oovs)))
            if summary != want:                                   # tolerated only where the oracle's own candidates tie
                gw, rw = summary.split(), want.split()
                k = next((j for j, (x, y) in enumerate(zip(gw, rw)) if x != y), min(len(gw), len(rw)))
                gaps = [abs(x[1] - y[1]) / max(1e-12, abs(x[1])) for step in tr[:k + 1] for x, y in zip(step, step[1:])
                        if np.isfinite(x[1]) and np.isfinite(y[1])]
                if not gaps or min(gaps) >= 2e-5:
                    raise SystemExit("bench.py: e2e row %d differs from the oracle at word %d with no near-tie" % (i, k))
                near += 1
            verified += 1
        if near > max(1, args.verify // 4):
            raise SystemExit("bench.py: %d of %d verified rows diverged at near-ties" % (near, args.verify))

    # ---------------- roofline: intervals inside one extra batch run with plain launches + CUDA events ----------
    # Only the first min_dec_steps (35) steps are averaged: no hypothesis may end before that (beam_search.py:146-149),
    # so every article is live in them and a launch does the full algorithmic work.
    dec.set_option("profile=35")
    dec.stage_upload(*batches[-1])
    dec.stage_run()
    dec.stage_sync()
    ps = dec.stats()
    dec.set_option("profile=0")
    n_prof = max(1.0, ps["prof_steps"])
    step_ms_vocab = ps["ms_vocab"] / n_prof
    step_ms_attn = ps["ms_attention"] / n_prof
    steps_done = dec.debug_fetch("steps_done", ARTICLES_PER_STEP, np.int32).astype(np.float64)
    # The headline durations: the two kernels replayed alone from a graph between two CUDA events on the library's
    # stream (no launch gaps, every article live), with a 256 MB overwrite between replays so that each starts from a
    # cold L2 like in a decode step; the flush-only graph is timed too and subtracted (Engine::measure).
    dec.set_option("measure=30")
    pm = dec.stats()
    n_rep = max(1.0, pm["prof_steps"])
    ms_attn, ms_vocab = pm["ms_attention"] / n_rep, pm["ms_vocab"] / n_rep
    pk = peaks()
    notes = profile_notes()
    R = ARTICLES_PER_STEP * BEAM
    lens = batches[-1][2].astype(np.float64)
    A = 2 * H
    # algorithmic bytes of the attention pass: F and h of every article once, coverage in, attn out
    attn_bytes = float(np.sum(2 * lens * A * 4 + 3 * BEAM * lens * 4))
    roof_attn = {"bound": "hbm", "kernel": "attention pass per decode step (attn_fused_kernel + attn_merge_kernel)",
                 "achieved": attn_bytes / (ms_attn * 1e-3) / 1e9 if ms_attn > 0 else None,
                 "peak": pk["hbm_gbs"], "unit": "GB/s", "ms_per_launch": ms_attn,
                 "ms_per_launch_in_profiled_step": step_ms_attn,
                 "timing": "graph replay x30 with an L2 flush between replays, flush time subtracted; the in-step figure is "
                           "the interval between two events around the plain launches of the first 35 steps (includes "
                           "launch gaps, and the feature re-packing kernel when the split decoder path runs)",
                 "algorithmic_bytes_per_launch": attn_bytes,
                 "traffic": notes.get("attention", {}).get("dram_bytes_per_launch"),
                 "traffic_source": notes.get("attention", {}).get("source"),
                 "bound_note": "the pass also needs 1.25 special-function operations per tanh: 65.5 M per step = 14 us "
                               "at 16 per clock and SM; the HBM bytes take 16 us at the measured peak"}
    if roof_attn["achieved"]:
        roof_attn["frac"] = roof_attn["achieved"] / pk["hbm_gbs"]
    try:
        tf32 = measure_tf32_peak(torch, dev)
    except Exception as ex:            # the peak measurement must never take the bench line down with it
        tf32 = {"tf32_tflops": None, "tf32_tflops_sustained": None, "how": "failed: %r" % (ex,)}
    flops = 2.0 * R * H * V
    roof_vocab = {"bound": "tensor", "kernel": "vocab_tc_kernel (tcgen05 kind::tf32) per decode step",
                  "achieved": flops / (ms_vocab * 1e-3) / 1e12 if ms_vocab > 0 else None,
                  "peak": tf32["tf32_tflops_sustained"], "peak_burst": tf32["tf32_tflops"], "peak_source": tf32["how"],
                  "unit": "TFLOP/s", "ms_per_launch": ms_vocab, "ms_per_launch_in_profiled_step": step_ms_vocab,
                  "bf16_peak_for_context": pk.get("bf16_tflops_sustained"),
                  "tensor_pipe_active_pct_ncu": notes.get("vocab_tc", {}).get("tensor_pipe_active_pct"),
                  "traffic": notes.get("vocab_tc", {}).get("dram_bytes_per_launch"),
                  "traffic_source": notes.get("vocab_tc", {}).get("source")}
    if roof_vocab["achieved"] and tf32["tf32_tflops_sustained"]:
        roof_vocab["frac"] = roof_vocab["achieved"] / tf32["tf32_tflops_sustained"]
    dominant = dict(roof_attn if ms_attn >= ms_vocab else roof_vocab, peak_source=pk.get("source"))

    # ---------------- config 2: encoder only (BASELINE.json configs[1]) ----------------
    enc_ms, rec_ms, feat_ms = ps["ms_encoder"], ps["ms_enc_recurrence"], ps["ms_enc_features"]
    n_cl = 2 * ((ARTICLES_PER_STEP + 9) // 10)                    # clusters of the recurrence: 2 directions x groups of 10 articles
    dsmem_bytes = float(lens.max()) * n_cl * 8 * 8 * 32 * 12 * 4  # per time step every CTA sends its 32 x 12 floats to 8 CTAs
    config2 = {"workload": "BASELINE.json configs[1]: encoder-only BiLSTM forward, 64 articles x 400 tokens",
               "encoder_ms": enc_ms, "recurrence_ms": rec_ms, "us_per_time_step": 1e3 * rec_ms / max(1.0, float(lens.max())),
               "dsmem_gbs": dsmem_bytes / (rec_ms * 1e-3) / 1e9 if rec_ms > 0 else None,
               "features_ms": feat_ms,
               "features_tflops_fp32_equivalent": 2.0 * float(lens.sum()) * A * A / (feat_ms * 1e-3) / 1e12 if feat_ms > 0 else None,
               "articles_per_s": ARTICLES_PER_STEP / (enc_ms * 1e-3) if enc_ms > 0 else None,
               "note": "input projection is a per-word table (no GEMM per batch); recurrence = one cluster kernel, "
                       "latency / DSMEM bound; features = 3xTF32 tcgen05 GEMM incl. the hi/lo split"}
    dec.close()

    # ---------------- config 4: beam 8, 120 steps, 200 OOVs per article (BASELINE.json configs[3]) ----------------
    c4_steps = max(2, min(4, args.steps))
    flags4 = synth.hyperparams_string(paths, beam_size=8, max_enc_steps=T_ENC, max_dec_steps=120, vocab_size=V,
                                      extra="--articles_per_batch=%d --flush_ms=-1" % ARTICLES_PER_STEP)
    dec4 = native.Decoder(flags4, device=local_rank)
    arts4 = synth.make_article_ids(ARTICLES_PER_STEP * 2, vsize=V, seed=4040 + rank, fixed_oovs=200)
    b4 = [native.pack_articles(arts4[i * ARTICLES_PER_STEP:(i + 1) * ARTICLES_PER_STEP], T_ENC) for i in range(2)]
    n4 = max(1, dec4.flag("slots"))
    str4 = []
    for sl in range(n4):
        dec4.stage_slot(sl)
        str4.append(torch.cuda.ExternalStream(dec4.stream_ptr(), device=dev))
        for _ in range(2):                                        # plain launches, then the graph capture
            dec4.stage_upload(*b4[sl % 2])
            dec4.stage_run()
    for sl in range(n4):
        dec4.stage_slot(sl)
        dec4.stage_sync()
    barrier()
    torch.cuda.synchronize()
    f0 = torch.cuda.Event(enable_timing=True)
    f0.record(str4[0])
    for i in range(c4_steps * n4):
        dec4.stage_slot(i % n4)
        dec4.stage_run()
    ends4 = []
    for sl in range(n4):
        f1 = torch.cuda.Event(enable_timing=True)
        f1.record(str4[sl])
        ends4.append(f1)
    for sl in range(n4):
        dec4.stage_slot(sl)
        dec4.stage_sync()
    torch.cuda.synchronize()
    ms4 = max_over_ranks(max(f0.elapsed_time(f1) for f1 in ends4))
    config4 = {"workload": "BASELINE.json configs[3]: beam 8, max_dec_steps 120, 200 distinct OOVs per article, 64 articles "
                           "per device batch", "articles_per_s": ARTICLES_PER_STEP * c4_steps * n4 * world / (ms4 * 1e-3),
               "ms_per_batch": ms4 / (c4_steps * n4), "batches_timed_per_gpu": c4_steps * n4, "n_gpus": world}
    dec4.close()

    out = {
        "metric": METRIC, "value": value, "unit": "articles/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(world),
        "e2e": {"value": e2e_value, "unit": "articles/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": "pgen_submit/pgen_poll with article byte strings"},
        "verified_rows": verified,
        "gpu_launches": launches, "clocks": clocks, "roofline": dominant,
        "roofline_other": roof_vocab if dominant.get("kernel") == roof_attn["kernel"] else roof_attn,
        "config2": config2, "config4": config4,
        "selection_guard": {"rows_widened": ps["selection_widened"], "rows_dense": ps["selection_fallbacks"]},
        "stage_ms_last_batch": {"h2d": ps["ms_h2d"], "encoder": ps["ms_encoder"], "decoder": ps["ms_decoder"],
                                "d2h": ps["ms_d2h"], "note": "the batch that carried the event pairs: its first %d decode steps were plain launches, the rest graph replays" % int(ps["prof_steps"])},
    }
    if rank == 0:
        try:
            threads = use_all_host_threads()
            sample_arts = synth.make_article_ids(args.cpu_articles, vsize=V, seed=99, full_len_only=False)
            secs = cpu_arm(paths, sample_arts, "literal")
            secs_d = cpu_arm(paths, sample_arts[:max(2, args.cpu_articles // 2)], "dedup")
            out["cpu_baseline"] = {"value": args.cpu_articles / secs, "unit": "articles/s", "cores": threads, "kind": "port",
                                   "dedup_value": max(2, args.cpu_articles // 2) / secs_d,
                                   "sample": "%d articles of the same synthetic workload, %.1f s; %s" % (args.cpu_articles, secs, CPU_KIND_NOTE)}
        except Exception as ex:      # the CPU baseline must never take the GPU number down with it
            out["cpu_baseline"] = {"value": None, "unit": "articles/s", "cores": os.cpu_count(), "kind": "port",
                                   "sample": "failed: %r" % (ex,)}
        print(json.dumps(out))
    if world > 1:
        sharding.wait_for_rank0(dist, "pgen_bench_cpu_leg", done=True)   # the other ranks sleep through rank 0's CPU leg
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
