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
