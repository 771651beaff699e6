"""Kernel-level parity through the C ABI: the attention pass on device buffers against the oracle."""
import numpy as np
import pytest

import pgen_oracle as O
from conftest import small_flags
from textsummarization_on_flink_b200 import native, synth

pytestmark = pytest.mark.gpu


def make_attention_case(torch, n, K, T, A, seed, lens=None):
    g = np.random.default_rng(seed)
    lens = np.asarray(lens if lens is not None else g.integers(1, T + 1, size=n), np.int32)
    enc = (g.standard_normal((n, T, A)) * 0.3).astype(np.float32)
    feat = (g.standard_normal((n, T, A)) * 0.8).astype(np.float32)
    dfeat = g.standard_normal((n * K, A)).astype(np.float32)
    cov = g.random((n * K, T)).astype(np.float32)
    host = dict(lens=lens, enc=enc, feat=feat, dfeat=dfeat, cov=cov)
    dev = {k: torch.from_numpy(v).cuda() for k, v in host.items()}
    return host, dev


def oracle_attention_from_features(W, host, K, use_cov):
    """oracle.attention with the decoder features given directly (bypassing the W_s linear layer)."""
    n, T, A = host["enc"].shape
    attn = np.zeros((n * K, T), np.float32)
    ctx = np.zeros((n * K, A), np.float32)
    v, wc = W[O.N_V], W[O.N_WC].reshape(-1)
    for b in range(n):
        L = int(host["lens"][b])
        for k in range(K):
            r = b * K + k
            x = host["feat"][b, :L] + host["dfeat"][r][None, :]
            if use_cov:
                x = x + host["cov"][r, :L, None] * wc[None, :]
            e = (v * np.tanh(x, dtype=np.float32)).sum(axis=1, dtype=np.float32)
            a = O.softmax(e[None])[0]
            a = a / a.sum(dtype=np.float32)
            attn[r, :L] = a
            ctx[r] = (a[:, None] * host["enc"][b, :L]).sum(axis=0, dtype=np.float32)
    return attn, ctx


@pytest.mark.parametrize("beam,coverage", [(4, 1), (4, 0), (8, 1), (2, 1)])
def test_attention_kernels_match_oracle(small_cfg, small_model_dir, beam, coverage):
    import torch
    T = 24
    flags = small_flags(small_model_dir, small_cfg, beam=beam, coverage=coverage, max_enc=T)
    A = 2 * small_cfg["H"]
    n = 7
    host, dev = make_attention_case(torch, n, beam, T, A, seed=beam)
    attn = torch.zeros(n * beam, T, device="cuda")
    ctx = torch.zeros(n * beam, A, device="cuda")
    with native.Decoder(flags) as dec:
        dec.attention(n, dev["feat"].data_ptr(), dev["enc"].data_ptr(), dev["lens"].data_ptr(), dev["dfeat"].data_ptr(),
                      dev["cov"].data_ptr(), attn.data_ptr(), ctx.data_ptr())
    ra, rc = oracle_attention_from_features(small_cfg["W"], host, beam, bool(coverage))
    ga, gc = attn.cpu().numpy(), ctx.cpu().numpy()
    for b in range(n):
        L = int(host["lens"][b])
        np.testing.assert_allclose(ga[b * beam:(b + 1) * beam, :L], ra[b * beam:(b + 1) * beam, :L], atol=5e-6, rtol=1e-4)
    np.testing.assert_allclose(gc, rc, atol=5e-6, rtol=1e-4)


@pytest.mark.parametrize("beam", [4, 8])
@pytest.mark.parametrize("impl", ["fused", "split"])
def test_full_size_attention_matches_oracle(model_dir, beam, impl):
    """The variants the decode loop runs at the benchmark's flags (hidden 256 -> A = 512, max_enc_steps 400, beam 4
    and 8): the one-launch kernel of csrc/attn_fused.cu and the attn_energy_pair / attn_context pair, against the
    oracle.  Ragged lengths cover full, one-position, exactly-one-segment and segment-plus-one articles."""
    import torch
    from textsummarization_on_flink_b200 import weights_io
    T, A = 400, 512
    flags = synth.hyperparams_string(model_dir, beam_size=beam, extra="--articles_per_batch=16 --slots=1")
    W = weights_io.load_weights(model_dir["weights"])
    lens = [400, 400, 1, 32, 33, 399, 64, 200, 17, 400, 385]
    n = len(lens)
    host, dev = make_attention_case(torch, n, beam, T, A, seed=10 + beam, lens=lens)
    attn = torch.full((n * beam, T), 7.0, device="cuda")             # poison: every entry must be written
    ctx = torch.full((n * beam, A), 7.0, device="cuda")
    with native.Decoder(flags) as dec:
        dec.set_option("attention=" + impl)
        for rep in range(2):                                         # twice: the segment counters must reset themselves
            dec.attention(n, dev["feat"].data_ptr(), dev["enc"].data_ptr(), dev["lens"].data_ptr(),
                          dev["dfeat"].data_ptr(), dev["cov"].data_ptr(), attn.data_ptr(), ctx.data_ptr())
    ra, rc = oracle_attention_from_features(W, host, beam, True)
    ga, gc = attn.cpu().numpy(), ctx.cpu().numpy()
    np.testing.assert_allclose(ga, ra, atol=2e-6, rtol=2e-4)         # zeros past the article included
    np.testing.assert_allclose(ga.sum(axis=1), 1.0, atol=2e-5)
    np.testing.assert_allclose(gc, rc, atol=5e-6, rtol=2e-4)


@pytest.mark.parametrize("use_tc", [0, 1])
def test_vocab_logits_against_float64(model_dir, use_tc):
    """out . w + v (model.py:236) for 256 decoder rows: the fp32 SIMT kernel within 1e-5 of a float64 product, the
    tcgen05 tf32 pass within 2e-3 (its scores only select candidates)."""
    import torch
    from textsummarization_on_flink_b200 import weights_io
    flags = synth.hyperparams_string(model_dir, extra="--articles_per_batch=64 --slots=1")
    W = weights_io.load_weights(model_dir["weights"])
    w = W["seq2seq/output_projection/w"].astype(np.float64)
    v = W["seq2seq/output_projection/v"].astype(np.float64)
    rows, H, V = 256, 256, w.shape[1]
    g = np.random.default_rng(5)
    out = (g.standard_normal((rows, H)) * 0.3).astype(np.float32)
    d_out = torch.from_numpy(out).cuda()
    d_log = torch.zeros(rows, V, device="cuda")
    with native.Decoder(flags) as dec:
        dec.vocab_logits(rows, d_out.data_ptr(), d_log.data_ptr(), bool(use_tc))
    exact = out.astype(np.float64) @ w + v
    err = np.abs(d_log.cpu().numpy() - exact).max()
    bound = 2e-3 * max(1.0, float(np.abs(exact).max())) if use_tc else 1e-5 * max(1.0, float(np.abs(exact).max()))
    assert err <= bound, (err, bound)
