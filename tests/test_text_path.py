"""Host text path of libpgen_cuda.so against the oracle (no GPU work, but pgen_create needs a device,
so these run with the GPU tests)."""
import numpy as np
import pytest

import pgen_oracle as O
from conftest import small_flags
from textsummarization_on_flink_b200 import native, synth

pytestmark = pytest.mark.gpu


def test_text_encode_and_summary_match_oracle(small_cfg, small_model_dir):
    V = small_cfg["V"]
    vocab = O.Vocab(small_model_dir["vocab_path"], V)
    rng = np.random.default_rng(0)
    arts = synth.make_articles(20, vsize=V, seed=4, max_len=40, oov_rate=0.3)
    arts += [b"", b" \t\n", b"article 0.", b"a\tb\nc\rd\x0be\x0cf  g", b"[STOP] [UNK] . , w7 w7 zz zz"]
    with native.Decoder(small_flags(small_model_dir, small_cfg)) as dec:
        for a in arts:
            ids, ext, n_oov = dec.text_encode(a)
            ex = O.make_example(a, vocab, 24)
            assert ids.tolist() == ex.enc_input and ext.tolist() == ex.enc_input_extend_vocab
            assert n_oov == len(ex.article_oovs)
            hi = V + n_oov
            toks = rng.integers(0, hi, size=30).tolist() if hi > 0 else []
            toks = [t if t != 3 else 4 for t in toks]
            for cut in (toks, toks[:7] + [3] + toks[7:], [3], [], [4, 4, 3]):
                want = O.format_summary(O.cut_at_stop(O.outputids2words(cut, vocab, ex.article_oovs)))
                assert dec.text_summary(cut) == want
            if n_oov == 0:
                with pytest.raises(native.PgenError):
                    dec.text_summary([V])


def test_flags_as_parsed_by_the_library(small_cfg, small_model_dir):
    with native.Decoder(small_flags(small_model_dir, small_cfg, beam=3, extra="--batch_size=2 --bogus=1")) as dec:
        assert dec.flag("beam_size") == 3 and dec.flag("batch_size") == 3      # run_summarization.py:312-313
        assert dec.flag("coverage") == 1 and dec.flag("pointer_gen") == 1 and dec.flag("min_dec_steps") == 8
        assert dec.flag("articles_per_batch") == 64 and dec.flag("nonexistent") == -1
