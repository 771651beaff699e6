"""Pins the oracle's text path (data.py / batcher.py / decode.py / run_summarization.py flags)."""
import pytest

import pgen_oracle as O


def _vocab(tmp_path, lines, max_size=0):
    p = tmp_path / "vocab"
    p.write_bytes(b"".join(l + b"\n" for l in lines))
    return O.Vocab(str(p), max_size)


def test_vocab_specials_and_order(tmp_path):
    v = _vocab(tmp_path, [b"the 100", b"cat 50", b". 40"])
    assert [v.id2word(i) for i in range(4)] == [b"[UNK]", b"[PAD]", b"[START]", b"[STOP]"]   # data.py:50-54
    assert v.word2id(b"the") == 4 and v.word2id(b"cat") == 5 and v.word2id(b".") == 6
    assert v.word2id(b"dog") == O.UNK_ID and v.size() == 7
    with pytest.raises(ValueError):
        v.id2word(7)


def test_vocab_max_size_counts_specials(tmp_path):
    v = _vocab(tmp_path, [b"a 3", b"b 2", b"c 1"], max_size=6)      # data.py:71-73
    assert v.size() == 6 and v.word2id(b"c") == O.UNK_ID


def test_vocab_bad_lines_skipped_and_errors(tmp_path):
    v = _vocab(tmp_path, [b"a 3", b"oops", b"x y z", b"b 2"])       # data.py:59-62
    assert v.size() == 6 and len(v.warnings) == 2
    with pytest.raises(ValueError):
        _vocab(tmp_path, [b"a 3", b"a 2"])                          # data.py:66-67
    with pytest.raises(ValueError):
        _vocab(tmp_path, [b"[UNK] 3"])                              # data.py:64-65
    with pytest.raises(ValueError):
        _vocab(tmp_path, [b"<s> 3"])


def test_article2ids_first_appearance_order(tmp_path):
    v = _vocab(tmp_path, [b"the 100", b"cat 50"])
    ids, oovs = O.article2ids(b"the zebra cat yak zebra the yak".split(), v)
    assert ids == [4, 6, 5, 7, 6, 4, 7] and oovs == [b"zebra", b"yak"]          # data.py:156-168
    assert O.outputids2words([4, 7, 3, 6], v, oovs) == [b"the", b"yak", b"[STOP]", b"zebra"]
    with pytest.raises(ValueError):
        O.outputids2words([8], v, oovs)


def test_make_example_truncates_and_splits_on_any_whitespace(tmp_path):
    v = _vocab(tmp_path, [b"the 100", b"cat 50"])
    ex = O.make_example(b"  the\tcat \n zebra  the ", v, 3)                      # batcher.py:51-56
    assert ex.enc_len == 3 and ex.enc_input == [4, 5, 0] and ex.enc_input_extend_vocab == [4, 5, 6]
    assert ex.article_oovs == [b"zebra"]
    ex0 = O.make_example(b"   ", v, 3)
    assert ex0.enc_len == 0 and ex0.enc_input == []


def test_tensorflowtest_article_shape(tmp_path):
    """TensorFlowTest.java:209 feeds "article 0." -> two tokens, the second an OOV."""
    v = _vocab(tmp_path, [b"article 10"])
    ex = O.make_example(b"article 0.", v, 400)
    assert ex.enc_input == [4, 0] and ex.enc_input_extend_vocab == [4, 5] and ex.article_oovs == [b"0."]


def test_format_summary():
    w = b"a b . c . d".split()
    assert O.format_summary(w) == b"a b .  c .  d"                               # decode.py:167-182
    assert O.format_summary([]) == b""
    assert O.format_summary([b".", b"."]) == b".  ."
    assert O.format_summary(b"< tag > .".split()) == b"< tag > ."                # make_html_safe is a no-op
    assert O.cut_at_stop([b"a", b"[STOP]", b"b", b"[STOP]"]) == [b"a"]
    assert O.cut_at_stop([b"a", b"b"]) == [b"a", b"b"]


def test_parse_hyperparams_like_the_reference_test():
    s = ("run_summarization.py --mode=decode --data_path=/x/0* --vocab_path=/v --log_root=/l "
         "--exp_name=pretrained_model_tf1.2.1 --batch_size=4 --max_enc_steps=400 --max_dec_steps=100 "
         "--coverage=1 --single_pass=1 --inference=1")                          # TensorFlowTest.java:40-53
    f = O.parse_hyperparams(s)
    assert f["mode"] == "decode" and f["coverage"] is True and f["single_pass"] is True
    assert f["beam_size"] == 4 and f["batch_size"] == 4 and f["min_dec_steps"] == 35
    assert f["vocab_size"] == 50000 and f["hidden_dim"] == 256 and f["emb_dim"] == 128 and f["pointer_gen"] is True


def test_parse_hyperparams_batch_size_dead_in_decode_and_unknown_ignored():
    f = O.parse_hyperparams("prog --mode=decode --batch_size=2 --beam_size=8 --articles_per_batch=64 --bogus")
    assert f["batch_size"] == 8                                                  # run_summarization.py:312-313
    f = O.parse_hyperparams("prog --mode=train --batch_size=2")
    assert f["batch_size"] == 2
    f = O.parse_hyperparams("prog --mode=decode --coverage=0 --nopointer_gen --coverage")
    assert f["pointer_gen"] is False and f["coverage"] is True
