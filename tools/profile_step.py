"""One warm batch, then one batch inside a cudaProfilerStart/Stop range (for ncu --profile-from-start off).
Usage: python tools/profile_step.py [articles_per_batch] [max_dec_steps] [extra flags, e.g. "--slots=1 --decoder=fused"]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
S = int(sys.argv[2]) if len(sys.argv) > 2 else 100
EXTRA = sys.argv[3] if len(sys.argv) > 3 else ""
paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
flags = synth.hyperparams_string(paths, max_dec_steps=S, min_dec_steps=min(35, S - 1),
                                 extra="--articles_per_batch=%d %s" % (B, EXTRA))
arts = synth.make_article_ids(2 * B, seed=1234)
with native.Decoder(flags) as dec:
    b0 = native.pack_articles(arts[:B], dec.max_enc_steps)
    b1 = native.pack_articles(arts[B:], dec.max_enc_steps)
    dec.stage_upload(*b0); dec.stage_run(); dec.stage_sync()
    dec.stage_upload(*b1); dec.stage_sync()
    torch.cuda.cudart().cudaProfilerStart()
    dec.stage_run(); dec.stage_sync()
    torch.cuda.cudart().cudaProfilerStop()
    st = dec.stats()
    print("encoder ms %.3f decoder ms %.3f steps %d launches %d" % (st["ms_encoder"], st["ms_decoder"], st["decode_steps"], st["kernel_launches"]))
