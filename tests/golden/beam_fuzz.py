"""Shared by make_reference_golden.py and tests/test_reference_golden.py: a deterministic stand-in for decode_onestep
(model.py:367-443) that makes beam search's bookkeeping the ONLY thing under test.  Log-probs are multiples of 0.25, so
sums and averages of different hypotheses tie exactly and the order of equal hypotheses (Python's stable sort,
beam_search.py:171-173), the [STOP] / min_dec_steps rule (:145-149) and the one-parent first step (:127-129) decide."""
import numpy as np

CASES = [dict(seed=s, K=[1, 2, 3, 4, 8][s % 5], V=6 + s % 7, n_oov=s % 3, min_dec=s % 5, max_dec=2 + s % 9) for s in range(160)]


def fake_step(case, step):
    """-> (ids [K, 2K] int32, logp [K, 2K] float32, descending per row like tf.nn.top_k's output)"""
    K, vext = case["K"], case["V"] + case["n_oov"]
    rng = np.random.RandomState((case["seed"] * 1000003 + step * 7919) & 0x7FFFFFFF)
    ids = np.zeros((K, 2 * K), np.int32)
    logp = np.zeros((K, 2 * K), np.float32)
    for i in range(K):
        ids[i] = rng.randint(0, vext, size=2 * K) if 2 * K > vext else rng.permutation(vext)[:2 * K]
        if rng.rand() < 0.5:
            ids[i, rng.randint(0, 2 * K)] = 3                       # [STOP] shows up often
        logp[i] = -0.25 * np.cumsum(rng.randint(0, 3, size=2 * K))  # descending, with equal neighbours
    return ids, logp
