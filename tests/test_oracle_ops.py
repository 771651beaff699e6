"""Pins the oracle's numerics op by op against independent formulations (the reference holds no
golden vectors -- SURVEY.md section 4 -- and TensorFlow cannot run here, so this is what stands in).
Tolerance: 2e-6 absolute on O(1) float32 values unless stated."""
import numpy as np
import pytest
import torch

import pgen_oracle as O
from textsummarization_on_flink_b200 import synth

ATOL = 2e-6


def _torch_lstm_from_tf(kernel, bias, E, H):
    """TF kernel [E+H,4H] with gate order i,j,f,o and forget_bias 1.0 -> torch.nn.LSTM (i,f,g,o)."""
    lstm = torch.nn.LSTM(E, H, batch_first=True)
    k = torch.from_numpy(kernel)
    i, j, f, o = torch.split(k, H, dim=1)
    perm = torch.cat([i, f, j, o], dim=1)                 # torch order: i, f, g(=j), o
    b = torch.from_numpy(bias)
    bi, bj, bf, bo = torch.split(b, H)
    with torch.no_grad():
        lstm.weight_ih_l0.copy_(perm[:E].T)
        lstm.weight_hh_l0.copy_(perm[E:].T)
        lstm.bias_ih_l0.copy_(torch.cat([bi, bf + 1.0, bj, bo]))
        lstm.bias_hh_l0.zero_()
    return lstm


def test_lstm_cell_matches_torch(small_cfg):
    H, E, W = small_cfg["H"], small_cfg["E"], small_cfg["W"]
    k, b = W[O.N_DEC_K], W[O.N_DEC_B]
    rng = np.random.default_rng(0)
    x = rng.standard_normal((5, E)).astype(np.float32)
    c = rng.standard_normal((5, H)).astype(np.float32)
    h = np.tanh(rng.standard_normal((5, H))).astype(np.float32)
    c2, h2 = O.lstm_cell(x, c, h, k, b)
    lstm = _torch_lstm_from_tf(k, b, E, H)
    with torch.no_grad():
        _, (ht, ct) = lstm(torch.from_numpy(x)[:, None, :], (torch.from_numpy(h)[None], torch.from_numpy(c)[None]))
    np.testing.assert_allclose(c2, ct[0].numpy(), atol=ATOL)
    np.testing.assert_allclose(h2, ht[0].numpy(), atol=ATOL)


@pytest.mark.parametrize("T", [1, 2, 9])
def test_encoder_matches_torch_bilstm(small_cfg, T):
    H, E, V, W = small_cfg["H"], small_cfg["E"], small_cfg["V"], small_cfg["W"]
    ids = np.random.default_rng(T).integers(0, V, size=T)
    enc, (c0, h0) = O.run_encoder(W, ids)
    x = torch.from_numpy(W[O.N_EMB][ids])[None]
    outs, fin = [], []
    for d in ("fw", "bw"):
        lstm = _torch_lstm_from_tf(W[O.N_ENC % (d, "kernel")], W[O.N_ENC % (d, "bias")], E, H)
        xi = x if d == "fw" else torch.flip(x, dims=[1])         # reverse_sequence over the own length
        with torch.no_grad():
            o, (hT, cT) = lstm(xi)
        outs.append(o[0] if d == "fw" else torch.flip(o[0], dims=[0]))
        fin.append((cT[0, 0], hT[0, 0]))
    ref = torch.cat(outs, dim=1).numpy()
    np.testing.assert_allclose(enc, ref, atol=ATOL)
    oc = torch.cat([fin[0][0], fin[1][0]]).numpy()
    oh = torch.cat([fin[0][1], fin[1][1]]).numpy()
    np.testing.assert_allclose(c0, np.maximum(oc @ W[O.N_RED % "w_reduce_c"] + W[O.N_RED % "bias_reduce_c"], 0), atol=ATOL)
    np.testing.assert_allclose(h0, np.maximum(oh @ W[O.N_RED % "w_reduce_h"] + W[O.N_RED % "bias_reduce_h"], 0), atol=ATOL)
    # backward half of position 0 is the backward final h (state after consuming token 0)
    np.testing.assert_allclose(enc[0, H:], fin[1][1].numpy(), atol=ATOL)


def test_softmax_and_topk_tie_rule():
    x = np.random.default_rng(1).standard_normal((3, 1000)).astype(np.float32) * 5
    np.testing.assert_allclose(O.softmax(x), torch.softmax(torch.from_numpy(x), -1).numpy(), atol=1e-7, rtol=1e-5)
    v = np.array([0.5, 0.9, 0.5, 0.9, 0.1, 0.9, 0.5] + [0.0] * 50, dtype=np.float32)
    vals, idx = O.top_k(v, 5)
    assert idx.tolist() == [1, 3, 5, 0, 2]                         # ties: lower index first (tf.nn.top_k)
    assert vals.tolist() == [np.float32(0.9)] * 3 + [0.5, 0.5]
    y = np.random.default_rng(2).standard_normal(5000).astype(np.float32)
    vals, idx = O.top_k(y, 8)
    tv, ti = torch.topk(torch.from_numpy(y), 8)
    assert idx.tolist() == ti.tolist() and np.array_equal(vals, tv.numpy())


def test_attention_matches_torch(small_cfg):
    H, W = small_cfg["H"], small_cfg["W"]
    A, R, T = 2 * H, 3, 11
    rng = np.random.default_rng(3)
    enc = rng.standard_normal((R, T, A)).astype(np.float32) * 0.3
    feat = O.encoder_features(W, enc)
    c = rng.standard_normal((R, H)).astype(np.float32)
    h = np.tanh(rng.standard_normal((R, H))).astype(np.float32)
    cov = rng.random((R, T)).astype(np.float32)
    mask = np.ones((R, T), np.float32)
    ctx, a, ncov = O.attention(W, feat, enc, mask, c, h, cov, True)
    tW = {k: torch.from_numpy(v).double() for k, v in W.items()}
    te, tc, th, tcov = (torch.from_numpy(z).double() for z in (enc, c, h, cov))
    tfeat = te @ tW[O.N_WH].reshape(A, A)
    d = torch.cat([tc, th], 1) @ tW[O.N_ATT_M] + tW[O.N_ATT_B]
    e = (tW[O.N_V] * torch.tanh(tfeat + d[:, None, :] + tcov[:, :, None] * tW[O.N_WC].reshape(1, 1, A))).sum(-1)
    ta = torch.softmax(e, -1)
    tctx = (ta[:, :, None] * te).sum(1)
    np.testing.assert_allclose(a, ta.numpy(), atol=5e-6)
    np.testing.assert_allclose(ctx, tctx.numpy(), atol=5e-6)
    np.testing.assert_allclose(ncov, (tcov + ta).numpy(), atol=5e-6)
    np.testing.assert_allclose(a.sum(1), 1.0, atol=1e-6)
    # no coverage -> coverage features absent, coverage output None
    ctx2, a2, nc2 = O.attention(W, feat, enc, mask, c, h, None, False)
    e2 = (tW[O.N_V] * torch.tanh(tfeat + d[:, None, :])).sum(-1)
    np.testing.assert_allclose(a2, torch.softmax(e2, -1).numpy(), atol=5e-6)
    assert nc2 is None


def test_final_dist_scatter_duplicates_accumulate(small_cfg):
    V, H, W = small_cfg["V"], small_cfg["H"], small_cfg["W"]
    rng = np.random.default_rng(4)
    R, T, n_oov = 2, 9, 2
    out = rng.standard_normal((R, H)).astype(np.float32)
    pg = rng.random((R, 1)).astype(np.float32)
    attn = O.softmax(rng.standard_normal((R, T)).astype(np.float32))
    ext = np.array([[5, 5, V, 7, V + 1, 5, V, 9, 7]] * R)
    hps = {"pointer_gen": True}
    fin = O.final_dist(W, hps, out, pg, attn, ext, n_oov)
    tv = torch.softmax(torch.from_numpy(out) @ torch.from_numpy(W[O.N_PROJ_W]) + torch.from_numpy(W[O.N_PROJ_V]), -1)
    ref = torch.zeros(R, V + n_oov)
    ref[:, :V] = torch.from_numpy(pg) * tv
    for r in range(R):
        ref[r].index_add_(0, torch.from_numpy(ext[r]), (1 - torch.from_numpy(pg[r])) * torch.from_numpy(attn[r]))
    np.testing.assert_allclose(fin, ref.numpy(), atol=1e-6)
    np.testing.assert_allclose(fin.sum(1), 1.0, atol=1e-5)         # a proper distribution over V+n_oov
    assert fin.shape == (R, V + n_oov)
    fin0 = O.final_dist(W, {"pointer_gen": False}, out, pg, attn, ext, n_oov)
    np.testing.assert_allclose(fin0, tv.numpy(), atol=1e-6)


def _articles(V, n, seed, max_len):
    return synth.make_article_ids(n, vsize=V, seed=seed, max_len=max_len, oov_rate=0.15)


@pytest.mark.parametrize("coverage", [1, 0])
@pytest.mark.parametrize("beam", [1, 3, 4])
def test_literal_equals_dedup(small_cfg, coverage, beam):
    """the first attention pass of step t>=1 has the same inputs as the second pass of
    step t-1, so carrying (context, attention) must reproduce the literal two-pass formulation."""
    V, W = small_cfg["V"], small_cfg["W"]
    hps = O.parse_hyperparams("p --mode=decode --beam_size=%d --coverage=%d --max_dec_steps=30 --min_dec_steps=8"
                              % (beam, coverage))
    for ids, ext, n_oov in _articles(V, 6, 11 + beam, 24):
        a = O.decode_ids(W, hps, ids, ext, n_oov, mode="literal")
        b = O.decode_ids(W, hps, ids, ext, n_oov, mode="dedup")
        assert a[0] == b[0]
        np.testing.assert_allclose(np.array(a[1]), np.array(b[1]), rtol=1e-5, atol=1e-6)


def test_beam_search_invariants(small_cfg):
    V, W = small_cfg["V"], small_cfg["W"]
    hps = O.parse_hyperparams("p --mode=decode --beam_size=4 --coverage=1 --max_dec_steps=25 --min_dec_steps=6")
    n_stop = 0
    for ids, ext, n_oov in _articles(V, 10, 5, 30):
        tr = []
        toks, lps, avg = O.decode_ids(W, hps, ids, ext, n_oov, mode="dedup", trace=tr)
        assert 1 <= len(toks) <= 25 and len(toks) == len(lps)
        assert all(0 <= t < V + n_oov for t in toks)
        if toks[-1] == O.STOP_ID:
            n_stop += 1
            assert len(toks) - 1 >= 6                          # [STOP] only accepted at steps >= min_dec_steps
        assert O.STOP_ID not in toks[:-1]
        assert abs(avg - sum(float(x) for x in lps) / (len(toks) + 1)) < 1e-12   # [START] counts in the length
        assert len(tr[0]) == 8 and all(len(s) == 32 for s in tr[1:])             # step 0 expands parent 0 only
        for s in tr:                                                              # sorted descending
            sc = [x[1] for x in s]
            assert sc == sorted(sc, reverse=True)
    assert n_stop >= 1


def test_zero_token_article_is_empty_summary(small_cfg):
    hps = O.parse_hyperparams("p --mode=decode")
    assert O.decode_ids(small_cfg["W"], hps, [], [], 0) == ([], [], 0.0)


def test_variable_inventory_matches_generator(small_cfg):
    assert O.variable_shapes(300, 32, 16) == synth.variable_shapes(300, 32, 16)
    assert {k: v.shape for k, v in small_cfg["W"].items()} == O.variable_shapes(300, 32, 16)
    n = sum(int(np.prod(s)) for s in O.variable_shapes(50000).values())
    assert 21_000_000 < n < 22_000_000                             # ~21.5 M parameters
