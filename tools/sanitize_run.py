"""A short run of every kernel of the decode path for compute-sanitizer (one tool per call, see profiles/):
small model (generic kernels, beam 3) and full-size model (persistent encoder, 3xTF32 features GEMM, fused decoder
cell, one-launch attention, tcgen05 vocabulary pass + guard, beam bookkeeping), plain launches and graph replay."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import native, synth, weights_io

full_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
# small model, generic kernels
V, H, E = 300, 32, 16
W = synth.make_weights("W1", vsize=V, hidden=H, emb=E, seed=7)
root = "/tmp/pgen_sanitize_small"
train = os.path.join(root, "log", "small", "train")
os.makedirs(train, exist_ok=True)
weights_io.save_weights(os.path.join(train, weights_io.DEFAULT_NAME), W)
synth.write_vocab_file(os.path.join(root, "vocab"), V)
flags = ("prog --mode=decode --vocab_path=%s/vocab --log_root=%s/log --exp_name=small --vocab_size=%d --hidden_dim=%d --emb_dim=%d "
         "--beam_size=3 --coverage=1 --max_enc_steps=24 --max_dec_steps=10 --min_dec_steps=3 --articles_per_batch=4 --slots=2" % (root, root, V, H, E))
arts = synth.make_article_ids(6, vsize=V, seed=3, max_len=24, oov_rate=0.15)
with native.Decoder(flags) as dec:
    tok, ln, lp, avg = dec.decode_ids(*native.pack_articles(arts, dec.max_enc_steps))
    for i, (a, e, k) in enumerate(arts):
        dec.submit(b"u%d" % i, synth.ids_to_article(e, V, salt=i), b"")
    rows = dec.poll(flush=True)
    print("small:", ln.tolist(), len(rows), "rows")
# full-size model, the benchmark's kernels; the second batch of a size replays the captured graph
paths = synth.make_model_dir("/tmp/pgen_bench_model", "W1")
for beam in (4, 8):
    fl = synth.hyperparams_string(paths, beam_size=beam, max_dec_steps=full_steps, min_dec_steps=2,
                                  extra="--articles_per_batch=8 --slots=1")
    arts = synth.make_article_ids(7, seed=21) + synth.make_article_ids(1, seed=22, max_len=5)
    with native.Decoder(fl) as dec:
        dec.set_option("graph_steps=2")
        packed = native.pack_articles(arts, dec.max_enc_steps)
        for rep in range(2):
            tok, ln, lp, avg = dec.decode_ids(*packed)
        dec.set_option("vocab_guard=dense")
        tok2, ln2, _, _ = dec.decode_ids(*packed)
        print("full beam %d:" % beam, ln.tolist(), "dense guard equal:", bool(np.array_equal(tok, tok2)), dec.stats()["kernel_launches"], "launches")
print("SANITIZE_RUN_OK")
