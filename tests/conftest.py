import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import pgen_pkg  # noqa: E402

pgen_pkg.load()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def small_cfg():
    """A tiny model (V=300, H=32, E=16) that the oracle decodes in milliseconds."""
    from textsummarization_on_flink_b200 import synth
    V, H, E = 300, 32, 16
    W = synth.make_weights("W1", vsize=V, hidden=H, emb=E, seed=7)
    W["seq2seq/output_projection/v"][3] = np.float32(4.0)
    return dict(V=V, H=H, E=E, W=W)


@pytest.fixture(scope="session")
def model_dir(tmp_path_factory):
    """Full-size synthetic model directory (vocab file + W1 weights archive), built once."""
    from textsummarization_on_flink_b200 import synth
    root = str(tmp_path_factory.mktemp("pgen_model"))
    return synth.make_model_dir(root, "W1")


@pytest.fixture(scope="session")
def small_model_dir(tmp_path_factory, small_cfg):
    """small_cfg written out as a vocab file + .pgenw archive for libpgen_cuda.so."""
    import os
    from textsummarization_on_flink_b200 import synth, weights_io
    root = str(tmp_path_factory.mktemp("pgen_small"))
    train = os.path.join(root, "log", "small", "train")
    os.makedirs(train)
    weights_io.save_weights(os.path.join(train, weights_io.DEFAULT_NAME), small_cfg["W"])
    synth.write_vocab_file(os.path.join(root, "vocab"), small_cfg["V"])
    return {"vocab_path": os.path.join(root, "vocab"), "log_root": os.path.join(root, "log"), "exp_name": "small"}


def small_flags(paths, cfg, beam=4, coverage=1, max_enc=24, max_dec=30, min_dec=8, extra=""):
    return ("prog --mode=decode --vocab_path=%s --log_root=%s --exp_name=%s --vocab_size=%d --hidden_dim=%d "
            "--emb_dim=%d --beam_size=%d --coverage=%d --max_enc_steps=%d --max_dec_steps=%d --min_dec_steps=%d "
            "--single_pass=1 --inference=1 %s" % (paths["vocab_path"], paths["log_root"], paths["exp_name"], cfg["V"],
                                                  cfg["H"], cfg["E"], beam, coverage, max_enc, max_dec, min_dec, extra))
