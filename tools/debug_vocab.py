import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth
paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
flags = synth.hyperparams_string(paths, extra="--articles_per_batch=64")
with native.Decoder(flags) as dec:
    for rows in (256, 100, 128):
        g = torch.Generator(device="cuda").manual_seed(rows)
        out = torch.randn(rows, 256, device="cuda", generator=g) * 0.5
        l_simt = torch.zeros(rows, 50000, device="cuda")
        l_tc = torch.zeros(rows, 50000, device="cuda")
        dec.vocab_logits(rows, out.data_ptr(), l_simt.data_ptr(), False)
        dec.vocab_logits(rows, out.data_ptr(), l_tc.data_ptr(), True)
        d = (l_simt - l_tc).abs()
        print("rows", rows, "simt std", l_simt.std().item(), "max abs diff", d.max().item(), "mean", d.mean().item(),
              "argmax agree", (l_simt.argmax(1) == l_tc.argmax(1)).float().mean().item())
        if d.max().item() > 0.1:
            bad = (d > 0.1).nonzero()
            print("bad count", len(bad), "first", bad[:5].tolist(), "rows", bad[:,0].unique()[:10].tolist(), "cols min/max", bad[:,1].min().item(), bad[:,1].max().item())
