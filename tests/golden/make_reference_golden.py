"""Golden vectors produced BY THE REFERENCE'S OWN PYTHON SOURCES, run in this container.

    python tests/golden/make_reference_golden.py            # needs /root/reference (read-only); writes ref_*.npz here

The reference's decode path (PG = /root/reference/src/main/python/pointer-generator) is Python 2 + TensorFlow 1.x.  Its
files are loaded from where they lie (never copied), with three in-memory adjustments so that CPython 3 can run them:
`print x` statements become calls, `xrange` is `range`, `import Queue` finds `queue` (no `/` of two integers occurs on the path).
One NumPy difference is undone: under NumPy >= 2 a Python float plus a float32 scalar stays float32, under the NumPy 1.x of
the reference's time it became float64 -- which is how `sum(self.log_probs)` (beam_search.py:72-79) got its float64 sums --
so decode_onestep's float32 log-probs reach run_beam_search as float64 copies of the same values.  `tensorflow` resolves to
oracle/tf1_shim.py, a NumPy restatement of the TF 1.11 ops these files call; `nltk` (third party, absent) resolves to a
two-function stand-in (whitespace tokens, sentences end at "."), `pyrouge` to an empty module.

What then runs is the reference's code, unmodified:
  PG/data.py            Vocab, article2ids, outputids2words, show_art_oovs ...
  PG/batcher.py         FlinkInferenceBatcher.next_batch (:565-585) -> FlinkExample / Example (:36-81) -> FlinkBatch / Batch
                        (:114-232); the Flink dataset it reads rows from is replaced by a list of (uuid, article, reference)
  PG/model.py           SummarizationModel.build_graph, run_encoder, decode_onestep (whole file, decode mode)
  PG/attention_decoder.py  attention_decoder, linear
  PG/beam_search.py     run_beam_search, Hypothesis, sort_hyps
  PG/decode.py          BeamSearchDecoder.__init__ / decode / decode_one_batch / write_for_flink (:42-185)
  PG/flink_writer.py    AbstractWriter (a subclass collects what write_result receives, :25-34)
  PG/util.py            get_config; load_ckpt is replaced: it assigns the synthetic weights to the graph's variables BY NAME
                        and fails on any variable the archive lacks or does not use -- which pins the variable names and
                        shapes of synth.variable_shapes / the .pgenw archive against the reference's own graph
The hyper-parameter tuple is built the way run_summarization.default_setup / inference_on_flink do (:294-331, :384-398:
batch_size = beam_size, the model gets max_dec_steps = 1).

Cases = the inputs of make_golden.py (SMALL: 12 articles on a 300-word model; FULL: 4 articles on the 50 000-word,
hidden-256 model), so that tests/test_reference_golden.py can state: the committed vectors the GPU tests compare with
are what the reference's code produces on these inputs.
"""
import builtins
import importlib.util
import os
import re
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import pgen_pkg  # noqa: E402

pgen_pkg.load()
import tf1_shim  # noqa: E402
from textsummarization_on_flink_b200 import synth  # noqa: E402

PG = "/root/reference/src/main/python/pointer-generator"
REF_FILES = ["data", "util", "flink_writer", "batcher", "attention_decoder", "model", "beam_search", "decode"]

# every flag run_summarization.py:47-87 defines, with its default
FLAG_DEFAULTS = dict(data_path="", vocab_path="", mode="train", num_steps=0, single_pass=False, inference=False, log_root="",
                     exp_name="", hidden_dim=256, emb_dim=128, batch_size=16, max_enc_steps=400, max_dec_steps=100, beam_size=4,
                     min_dec_steps=35, vocab_size=50000, lr=0.15, adagrad_init_acc=0.1, rand_unif_init_mag=0.02,
                     trunc_norm_init_std=1e-4, max_grad_norm=2.0, pointer_gen=True, coverage=False, cov_loss_wt=1.0,
                     convert_to_coverage_model=False, restore_best_model=False, debug=False)
HPARAM_LIST = ['mode', 'lr', 'adagrad_init_acc', 'rand_unif_init_mag', 'trunc_norm_init_std', 'max_grad_norm', 'hidden_dim',
               'emb_dim', 'batch_size', 'max_dec_steps', 'max_enc_steps', 'coverage', 'cov_loss_wt', 'pointer_gen', 'inference',
               'num_steps']                                            # run_summarization.py:319-321


def _py3(src):
    """Python 2 -> 3, in memory: print statements, nothing else changes the meaning of these files."""
    src = re.sub(r"^(\s*)print[ \t]+(?!\()(.+)$", r"\1print(\2)", src, flags=re.M)
    src = re.sub(r"^import Queue$", "import queue as Queue", src, flags=re.M)
    return src


def load_reference():
    """-> dict name -> module of the reference's files, executing under the TF shim"""
    if not os.path.isdir(PG):
        raise SystemExit("make_reference_golden.py runs only where /root/reference exists")
    mods = tf1_shim.build_modules()
    nltk = types.ModuleType("nltk")                                    # third party, absent: whitespace tokens, "." ends a sentence
    nltk.word_tokenize = lambda s: s.split()

    def sent_tokenize(s):
        out, cur = [], []
        for w in s.split():
            cur.append(w)
            if w == ".":
                out.append(" ".join(cur))
                cur = []
        if cur:
            out.append(" ".join(cur))
        return out
    nltk.sent_tokenize = sent_tokenize
    mods["nltk"] = nltk
    mods["pyrouge"] = types.ModuleType("pyrouge")
    sys.modules.update(mods)
    builtins.xrange = range
    out = {}
    for name in REF_FILES:
        path = os.path.join(PG, name + ".py")
        m = types.ModuleType(name)
        m.__file__ = path
        sys.modules[name] = m
        exec(compile(_py3(open(path).read()), path, "exec"), m.__dict__)
        out[name] = m
    out["tf"] = mods["tensorflow"]
    return out


class Collect(object):
    """what one reference decode run produced, row by row"""

    def __init__(self):
        self.rows, self.hyps, self.steps, self.batches = [], [], [], []


def run_reference(R, W, vocab_path, vsize, hidden, emb, beam, max_enc, max_dec, min_dec, coverage, rows, pointer_gen=True):
    """rows: list of (uuid, article, reference) strings -> Collect.  Everything between the row list and the collected
    (uuid, article, summary, reference) is the reference's code."""
    tf = R["tf"]
    tf.reset_default_graph()
    if not coverage:                                                   # a graph without coverage has no w_c (attention_decoder.py:70-72)
        W = {k: v for k, v in W.items() if not k.endswith("coverage/w_c")}
    if not pointer_gen:                                                # ... and the baseline model has no p_gen layer (:163-167)
        W = {k: v for k, v in W.items() if "calculate_pgen" not in k}
    FLAGS = tf.app.flags.FLAGS
    for k, v in FLAG_DEFAULTS.items():
        setattr(FLAGS, k, v)
    log_root = tempfile.mkdtemp(prefix="pgen_ref_log_")
    FLAGS.mode, FLAGS.single_pass, FLAGS.inference, FLAGS.data_path = "decode", True, True, "test"
    FLAGS.vocab_path, FLAGS.vocab_size, FLAGS.log_root = vocab_path, vsize, log_root
    FLAGS.hidden_dim, FLAGS.emb_dim, FLAGS.beam_size = hidden, emb, beam
    FLAGS.max_enc_steps, FLAGS.max_dec_steps, FLAGS.min_dec_steps, FLAGS.coverage = max_enc, max_dec, min_dec, bool(coverage)
    FLAGS.pointer_gen = bool(pointer_gen)
    # default_setup (run_summarization.py:294-331)
    vocab = R["data"].Vocab(FLAGS.vocab_path, FLAGS.vocab_size)
    FLAGS.batch_size = FLAGS.beam_size
    import collections
    hps_dict = {k: getattr(FLAGS, k) for k in HPARAM_LIST}
    hps = collections.namedtuple("HParams", hps_dict.keys())(**hps_dict)
    col = Collect()

    # the Flink dataset of FlinkInferenceBatcher: `self._sess.run(self._next_batch)` yields one row, then OutOfRange
    class RowSource(object):
        def __init__(self):
            self.i = 0

        def run(self, _fetch):
            if self.i >= len(rows):
                raise tf.errors.OutOfRangeError()
            u, a, r = rows[self.i]
            self.i += 1
            return (np.array([u], object), np.array([a], object), np.array([r], object))
    batcher = object.__new__(R["batcher"].FlinkInferenceBatcher)      # its __init__ only builds that dataset (batcher.py:540-552)
    batcher._sess, batcher._next_batch, batcher._vocab, batcher._hps = RowSource(), None, vocab, hps
    inner_next = batcher.next_batch

    def next_batch():
        b = inner_next()
        if b is not None:
            col.batches.append(dict(enc_batch=b.enc_batch.copy(), enc_lens=b.enc_lens.copy(), enc_padding_mask=b.enc_padding_mask.copy(),
                                    enc_batch_extend_vocab=b.enc_batch_extend_vocab.copy() if pointer_gen else b.enc_batch.copy(),
                                    max_art_oovs=int(b.max_art_oovs) if pointer_gen else 0,
                                    art_oovs=list(b.art_oovs[0]) if pointer_gen else []))
        return b
    batcher.next_batch = next_batch

    class Writer(R["flink_writer"].AbstractWriter):
        def write_result(self, sess, uuid, article, abstract, reference):
            col.rows.append((uuid, article, abstract, reference))

    def load_ckpt(saver, sess, ckpt_dir="train"):
        used = set()
        for v in tf.global_variables():
            if v.var_name == "global_step":
                continue
            if v.var_name not in W:
                raise KeyError("the reference's graph has variable %r, the weight archive does not" % v.var_name)
            v.assign_value(W[v.var_name])
            used.add(v.var_name)
        extra = sorted(set(W) - used)
        if extra:
            raise KeyError("the weight archive has variables the reference's graph never creates: %r" % extra)
        return os.path.join(FLAGS.log_root, ckpt_dir, "model.ckpt-0")
    R["util"].load_ckpt = load_ckpt

    decode_model_hps = hps._replace(max_dec_steps=1)                   # run_summarization.py:391
    model = R["model"].SummarizationModel(decode_model_hps, vocab)
    real_onestep = model.decode_onestep

    def onestep(*a, **k):
        out = real_onestep(*a, **k)
        col.steps[-1].append((np.array(out[0]), np.array(out[1])))
        # NumPy 1.x semantics for beam_search.py:72-79: `sum(self.log_probs)` starts from the Python float 0.0, and the
        # NumPy the reference ran on added a float32 scalar to it in float64; NumPy >= 2 (NEP 50) would stay in float32.
        # The same float32 values are therefore handed on as float64, which gives the sums the reference computed.
        return (out[0], np.asarray(out[1], np.float64)) + tuple(out[2:])
    model.decode_onestep = onestep
    real_search = R["beam_search"].run_beam_search

    def search(sess, mdl, voc, batch):
        col.steps.append([])
        h = real_search(sess, mdl, voc, batch)
        col.hyps.append(h)
        return h
    R["beam_search"].run_beam_search = search
    try:
        decoder = R["decode"].BeamSearchDecoder(model, batcher, vocab, writer=Writer())
        decoder.decode(withRouge=False)
    finally:
        R["beam_search"].run_beam_search = real_search
    col.variables = {v.var_name: tuple(v.value.shape) for v in tf.global_variables() if v.var_name != "global_step"}
    return col


def pack(col, S, K2):
    n = len(col.hyps)
    out = dict(tokens=np.zeros((n, S), np.int32), lens=np.zeros((n,), np.int32), logprobs=np.zeros((n, S), np.float32),
               avg=np.zeros((n,), np.float64), n_steps=np.zeros((n,), np.int32),
               step_ids=np.zeros((n, S, K2 // 2, K2), np.int32), step_logp=np.zeros((n, S, K2 // 2, K2), np.float32))
    for b, h in enumerate(col.hyps):
        t = [int(x) for x in h.tokens[1:]]                             # decode.py:108
        out["lens"][b] = len(t)
        out["tokens"][b, :len(t)] = t
        out["logprobs"][b, :len(t)] = np.array(h.log_probs[1:], np.float32)
        out["avg"][b] = h.avg_log_prob
        out["n_steps"][b] = len(col.steps[b])
        for s, (ids, lp) in enumerate(col.steps[b]):
            out["step_ids"][b, s] = ids
            out["step_logp"][b, s] = lp
    return out


def small_case(R):
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    c = mg.SMALL
    tmp = tempfile.mkdtemp(prefix="pgen_ref_small_")
    vpath = os.path.join(tmp, "vocab")
    synth.write_vocab_file(vpath, c["V"])
    arts = synth.make_article_ids(c["n"], vsize=c["V"], seed=c["art_seed"], max_len=c["max_len"], oov_rate=c["oov_rate"])
    rows = [("uuid-%d" % i, synth.ids_to_article(ext, c["V"], salt=i).decode(), "reference %d . second sentence ." % i)
            for i, (_, ext, _) in enumerate(arts)]
    col = run_reference(R, mg.small_weights(), vpath, c["V"], c["H"], c["E"], 4, 24, 30, 8, 1, rows)
    return mg, arts, rows, col


def full_case(R, mg):
    c = mg.FULL
    tmp = tempfile.mkdtemp(prefix="pgen_ref_full_")
    vpath = os.path.join(tmp, "vocab")
    synth.write_vocab_file(vpath, 50000)
    arts = synth.make_article_ids(c["n"], seed=c["art_seed"])
    rows = [("uuid-%d" % i, synth.ids_to_article(ext, 50000, salt=i).decode(), "reference %d ." % i) for i, (_, ext, _) in enumerate(arts)]
    col = run_reference(R, synth.make_weights("W1"), vpath, 50000, 256, 128, 4, 400, 100, 35, 1, rows)
    return arts, rows, col


# More of the flag space and of the text path on the 300-word model: (beam_size, coverage, max_dec_steps, min_dec_steps).
VARIANTS = [(8, 1, 20, 5), (3, 0, 16, 4), (1, 1, 12, 3), (2, 0, 12, 12)]
# article texts chosen for the text path: truncation at max_enc_steps, the special tokens spelled out in an article
# ([UNK] maps to the UNK id and therefore becomes an in-article OOV, data.py:152-158), repeated OOVs, one-word and
# all-OOV articles, runs of blanks / tabs / newlines (str.split), punctuation ranks 4 and 5
EDGE_ARTICLES = [
    " ".join("w%d" % (6 + (i * 7) % 200) for i in range(40)),
    "[STOP] [UNK] . , w7 w7 zz zz [PAD] [START] zz .",
    "w9",
    "qq",
    "aa bb aa cc bb aa . dd , aa",
    "a\tb\nc\rd\x0be\x0cf  g   w10  .",
    "w20 w21 . w22 w23 . w24 , w25 . . w26",
]


def variants_case(R, mg):
    c = mg.SMALL
    tmp = tempfile.mkdtemp(prefix="pgen_ref_var_")
    vpath = os.path.join(tmp, "vocab")
    synth.write_vocab_file(vpath, c["V"])
    arts = synth.make_article_ids(5, vsize=c["V"], seed=321, max_len=c["max_len"], oov_rate=0.25)
    rows = [("v-%d" % i, synth.ids_to_article(ext, c["V"], salt=i).decode(), "ref %d ." % i) for i, (_, ext, _) in enumerate(arts)]
    out = {}
    for beam, cov, max_dec, min_dec in VARIANTS:
        col = run_reference(R, mg.small_weights(), vpath, c["V"], c["H"], c["E"], beam, 24, max_dec, min_dec, cov, rows)
        tag = "b%dc%d_" % (beam, cov)
        pk = pack(col, max_dec, 2 * beam)
        for k in ("tokens", "lens", "logprobs", "avg", "n_steps", "step_ids", "step_logp"):
            out[tag + k] = pk[k]
        out[tag + "summary"] = np.array([r[2] for r in col.rows])
    # the baseline model (pointer_gen off: OOVs stay [UNK], no copy distribution) and an all-zero model, where every
    # vocabulary word ties and only the rules decide: tf.nn.top_k takes the lower id first, Python's sorted(...,
    # reverse=True) keeps equal hypotheses in their original order
    rows5 = rows
    col = run_reference(R, mg.small_weights(), vpath, c["V"], c["H"], c["E"], 4, 24, 14, 4, 1, rows5, pointer_gen=False)
    pk = pack(col, 14, 8)
    for k in ("tokens", "lens", "logprobs", "avg", "n_steps", "step_ids", "step_logp"):
        out["nopg_" + k] = pk[k]
    out["nopg_summary"] = np.array([r[2] for r in col.rows])
    zero = {k: np.zeros_like(v) for k, v in mg.small_weights().items()}
    col = run_reference(R, zero, vpath, c["V"], c["H"], c["E"], 4, 24, 8, 3, 1, rows5)
    pk = pack(col, 8, 8)
    for k in ("tokens", "lens", "logprobs", "avg", "n_steps", "step_ids", "step_logp"):
        out["zero_" + k] = pk[k]
    out["zero_summary"] = np.array([r[2] for r in col.rows])
    rows = [("e-%d" % i, a, "edge reference %d ." % i) for i, a in enumerate(EDGE_ARTICLES)]
    col = run_reference(R, mg.small_weights(), vpath, c["V"], c["H"], c["E"], 4, 24, 30, 8, 1, rows)
    pk = pack(col, 30, 8)
    for k in ("tokens", "lens", "logprobs", "avg"):
        out["edge_" + k] = pk[k]
    out["edge_article"] = np.array(EDGE_ARTICLES)
    out["edge_summary"] = np.array([r[2] for r in col.rows])
    out["edge_enc_lens"] = np.array([int(b["enc_lens"][0]) for b in col.batches], np.int32)
    out["edge_enc"] = np.array([" ".join(map(str, b["enc_batch"][0][:int(b["enc_lens"][0])])) for b in col.batches])
    out["edge_ext"] = np.array([" ".join(map(str, b["enc_batch_extend_vocab"][0][:int(b["enc_lens"][0])])) for b in col.batches])
    out["edge_oovs"] = np.array(["\t".join(b["art_oovs"]) for b in col.batches])
    return out


# Vocabulary files for data.py:40-75 (ASCII, "\n" line ends only: under Python 3 the reference's open(..., 'r') would also
# break lines at a lone "\r", which Python 2 did not).  (contents, max_size)
VOCAB_FILES = [
    ("a 10\nb 9\nc 8\nd 7\ne 6\nf 5\n", 0),
    ("a 10\nb 9\nc 8\nd 7\ne 6\nf 5\n", 8),
    ("a 10\nb 9\nc 8\n", 3),                                     # fewer than the 4 specials: one word is still read
    ("a 10\nlonely\nb 9 extra\n\n   \nc 8\n", 0),               # 1 field, 3 fields, empty and blank lines are skipped
    ("a\t10\n  b   9  \nc 8", 0),                                 # tabs, runs of blanks, no newline at the end
    ("a 10\nb 9\na 8\n", 0),                                      # duplicate
    ("a 10\nb 9\na 8\n", 6),                                      # ... not reached: reading stops at max_size
    ("a 10\n[UNK] 9\n", 0),
    ("a 10\n<s> 9\n", 0),
    ("a 10\n</s> 9\nb 1\n", 0),
    ("a 10\n[STOP]\nb 1\n", 0),                                    # a special token on a malformed line is only skipped
    ("A 10\na 9\n. 8\n, 7\n", 0),                                 # case matters
]


def vocab_case(R):
    out = []
    tmp = tempfile.mkdtemp(prefix="pgen_ref_vocab_")
    for i, (text, max_size) in enumerate(VOCAB_FILES):
        path = os.path.join(tmp, "vocab%d" % i)
        with open(path, "w", newline="") as f:
            f.write(text)
        try:
            v = R["data"].Vocab(path, max_size)
            words = [v.id2word(j) for j in range(v.size())]
            out.append("ok\t%d\t%s" % (v.size(), " ".join(words)))
        except Exception as ex:                                          # data.py:65,67 raise plain Exception
            out.append("error\t%s" % ex)
    return out


def beam_case(R):
    """run_beam_search / Hypothesis / sort_hyps of the reference (beam_search.py, unmodified) on the deterministic stand-in
    for decode_onestep of beam_fuzz.py: exact ties everywhere."""
    import beam_fuzz
    tf = R["tf"]
    FLAGS = tf.app.flags.FLAGS
    LST = tf.contrib.rnn.LSTMStateTuple
    out_tok, out_lp, out_avg = [], [], []
    for case in beam_fuzz.CASES:
        FLAGS.beam_size, FLAGS.min_dec_steps, FLAGS.max_dec_steps = case["K"], case["min_dec"], case["max_dec"]
        FLAGS.pointer_gen, FLAGS.coverage = True, True
        K = case["K"]

        class Vocab(object):
            def size(self): return case["V"]
            def word2id(self, w): return {"[UNK]": 0, "[PAD]": 1, "[START]": 2, "[STOP]": 3}[w]

        class Model(object):
            def __init__(self): self.step = 0
            def run_encoder(self, sess, batch): return np.zeros((K, 5, 2), np.float32), LST(np.zeros(1, np.float32), np.zeros(1, np.float32))
            def decode_onestep(self, sess, batch, latest_tokens, enc_states, dec_init_states, prev_coverage):
                assert all(0 <= t < case["V"] for t in latest_tokens)            # OOV ids were mapped to [UNK] (:112-113)
                ids, logp = beam_fuzz.fake_step(case, self.step)
                self.step += 1
                logp = logp.astype(np.float64)                                   # NumPy 1.x promotion, see run_reference
                return (ids, logp, [dec_init_states[0]] * K, [[0.2] * 5] * K, [0.5] * K, [np.zeros(5)] * K)
        batch = types.SimpleNamespace(enc_batch=np.zeros((K, 5), np.int32))
        h = R["beam_search"].run_beam_search(None, Model(), Vocab(), batch)
        out_tok.append([int(t) for t in h.tokens])
        out_lp.append([float(x) for x in h.log_probs])
        out_avg.append(float(h.avg_log_prob))
    n = max(len(t) for t in out_tok)
    tok = np.full((len(out_tok), n), -1, np.int32)
    lp = np.zeros((len(out_tok), n), np.float64)
    for i, (t, l) in enumerate(zip(out_tok, out_lp)):
        tok[i, :len(t)] = t
        lp[i, :len(l)] = l
    return dict(beam_tokens=tok, beam_logp=lp, beam_avg=np.array(out_avg, np.float64))


def flag_table():
    """The tf.app.flags.DEFINE_* statements of run_summarization.py:47-87, EXECUTED (not re-typed) against a recorder."""
    src = open(os.path.join(PG, "run_summarization.py")).read()
    stmts = re.findall(r"^tf\.app\.flags\.DEFINE_\w+\(.*\)$", src, flags=re.M)
    table = []

    class Rec(object):
        def __getattr__(self, kind):
            return lambda name, default, doc: table.append((kind.replace("DEFINE_", ""), name, default))
    tf = types.SimpleNamespace(app=types.SimpleNamespace(flags=Rec()))
    for st in stmts:
        exec(st, {"tf": tf})
    return table


def main():
    R = load_reference()
    mg, arts, rows, col = small_case(R)
    out = pack(col, 30, 8)
    # the row-level path: what the writer received (uuid, article, summary, reference), and the batch arrays of row 0..n
    out["row_uuid"] = np.array([r[0] for r in col.rows])
    out["row_article"] = np.array([r[1] for r in col.rows])
    out["row_summary"] = np.array([r[2] for r in col.rows])
    out["row_reference"] = np.array([r[3] for r in col.rows])
    out["in_reference"] = np.array([r[2] for r in rows])
    T = 24
    eb = np.zeros((len(rows), T), np.int32)
    ee = np.zeros((len(rows), T), np.int32)
    for i, b in enumerate(col.batches):
        assert (b["enc_batch"] == b["enc_batch"][0]).all() and b["enc_batch"].shape[0] == 4      # one example repeated beam_size times
        n = int(b["enc_lens"][0])
        eb[i, :n] = b["enc_batch"][0]
        ee[i, :n] = b["enc_batch_extend_vocab"][0]
        assert b["enc_padding_mask"][0].sum() == n
    out["enc_batch"], out["enc_batch_extend_vocab"] = eb, ee
    out["enc_lens"] = np.array([int(b["enc_lens"][0]) for b in col.batches], np.int32)
    out["n_oovs"] = np.array([b["max_art_oovs"] for b in col.batches], np.int32)
    out["art_oovs"] = np.array(["\t".join(b["art_oovs"]) for b in col.batches])
    out["variable_names"] = np.array(sorted(col.variables))
    out["variable_shapes"] = np.array([",".join(map(str, col.variables[k])) for k in sorted(col.variables)])
    np.savez_compressed(os.path.join(HERE, "ref_small.npz"), **out)
    print("ref_small.npz: %d rows, summary lengths %s" % (len(col.rows), out["lens"].tolist()))
    var = variants_case(R, mg)
    var["vocab_file"] = np.array([t for t, _ in VOCAB_FILES])
    var["vocab_max"] = np.array([m for _, m in VOCAB_FILES], np.int64)
    var["vocab_outcome"] = np.array(vocab_case(R))
    var.update(beam_case(R))
    ft = flag_table()
    assert {n: d for _, n, d in ft} == FLAG_DEFAULTS, "FLAG_DEFAULTS above and run_summarization.py differ"
    var["flag_kind"] = np.array([k for k, _, _ in ft])
    var["flag_name"] = np.array([n for _, n, _ in ft])
    var["flag_default"] = np.array([repr(d) for _, _, d in ft])
    np.savez_compressed(os.path.join(HERE, "ref_variants.npz"), **var)
    print("ref_variants.npz:", {k: var[k].tolist() for k in var if k.endswith("lens")})
    if "--small-only" not in sys.argv:
        arts, rows, col = full_case(R, mg)
        out = pack(col, 100, 8)
        out["row_summary"] = np.array([r[2] for r in col.rows])
        out["variable_names"] = np.array(sorted(col.variables))
        out["variable_shapes"] = np.array([",".join(map(str, col.variables[k])) for k in sorted(col.variables)])
        # BASELINE.json configs[0] says "random-init weights": W0 = the reference's own initialisers (synth.make_weights)
        arts0 = synth.make_article_ids(2, seed=4242)
        rows0 = [("w0-%d" % i, synth.ids_to_article(ext, 50000, salt=i).decode(), "reference %d ." % i) for i, (_, ext, _) in enumerate(arts0)]
        vpath = os.path.join(tempfile.mkdtemp(prefix="pgen_ref_w0_"), "vocab")
        synth.write_vocab_file(vpath, 50000)
        col0 = run_reference(R, synth.make_weights("W0"), vpath, 50000, 256, 128, 4, 400, 100, 35, 1, rows0)
        pk0 = pack(col0, 100, 8)
        for k in ("tokens", "lens", "logprobs", "avg"):
            out["w0_" + k] = pk0[k]
        np.savez_compressed(os.path.join(HERE, "ref_full.npz"), **out)
        print("ref_full.npz: %d rows, summary lengths %s" % (len(col.rows), out["lens"].tolist()))


if __name__ == "__main__":
    main()
