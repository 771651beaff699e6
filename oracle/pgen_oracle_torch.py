"""TEST INFRASTRUCTURE -- PyTorch-CPU restatement of the numeric core of the reference's decode path.

Second, independent implementation of what oracle/pgen_oracle.py restates in NumPy (same reference lines, cited per
function): the graph side (encoder, attention decoder, projection, pointer mixture) runs as torch CPU ops in a chosen
dtype, the host side (beam_search.py:82-168) is pgen_oracle.run_beam_search itself -- exactly the split the reference
has between `sess.run` and its Python beam search.  Two uses:

* float64: a whole-decode reference in (nearly) exact arithmetic, to put an error bar on the float32 NumPy oracle
  (tests/test_oracle_torch.py);
* float32 with all host threads: the CPU baseline BASELINE.md section 3 names (bench.py `cpu_baseline` and
  `--impl reference`; TensorFlow 1.x itself cannot run here).

PARITY UNPINNED by the reference (it holds no golden vectors and TF1 is unavailable): this module pins the oracle
against an independent formulation, not against TensorFlow.  Only tests/, __graft_entry__.smoke() and bench.py's CPU
legs may import it; the product never does.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import numpy as np
import torch

import pgen_oracle as O

_CACHE: Dict[tuple, Dict[str, torch.Tensor]] = {}


def use_all_threads() -> int:
    """torchrun exports OMP_NUM_THREADS=1; the CPU arm is meant to use every host core."""
    n = os.cpu_count() or 1
    torch.set_num_threads(n)
    return torch.get_num_threads()


def _weights(W, dtype) -> Dict[str, torch.Tensor]:
    key = (id(W), dtype)
    if key not in _CACHE:
        _CACHE.clear()
        _CACHE[key] = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dtype) for k, v in W.items()}
    return _CACHE[key]


class TorchGraph:
    """The functions run_beam_search calls into, in torch.  NumPy in, NumPy out (like sess.run)."""

    def __init__(self, W, dtype=torch.float32):
        self.W = W
        self.Wt = _weights(W, dtype)
        self.dt = dtype
        self.npdt = np.float64 if dtype == torch.float64 else np.float32

    def t(self, x) -> torch.Tensor:
        return torch.from_numpy(np.ascontiguousarray(x)).to(self.dt)

    # LSTMCell (model.py:90-91,138): gates i, j, f, o; forget_bias 1.0
    def lstm_cell(self, x, c, h, kernel, bias):
        z = torch.cat([x, h], dim=-1) @ kernel + bias
        i, j, f, o = torch.chunk(z, 4, dim=-1)
        c2 = torch.sigmoid(f + 1.0) * c + torch.sigmoid(i) * torch.tanh(j)
        return c2, torch.sigmoid(o) * torch.tanh(c2)

    def run_encoder(self, W, enc_ids):
        """model.py:210-221, 76-121, 347-364 for one article."""
        Wt = self.Wt
        T = len(enc_ids)
        emb = Wt[O.N_EMB][torch.as_tensor(np.asarray(enc_ids, dtype=np.int64))]
        H = Wt[O.N_RED % "bias_reduce_c"].shape[0]
        out = torch.zeros((T, 2 * H), dtype=self.dt)
        fin = {}
        for d, order in (("fw", range(T)), ("bw", range(T - 1, -1, -1))):
            k, b = Wt[O.N_ENC % (d, "kernel")], Wt[O.N_ENC % (d, "bias")]
            c = torch.zeros((H,), dtype=self.dt)
            h = torch.zeros((H,), dtype=self.dt)
            off = 0 if d == "fw" else H
            for t in order:
                c, h = self.lstm_cell(emb[t], c, h, k, b)
                out[t, off:off + H] = h
            fin[d] = (c, h)
        old_c = torch.cat([fin["fw"][0], fin["bw"][0]])
        old_h = torch.cat([fin["fw"][1], fin["bw"][1]])
        c0 = torch.relu(old_c @ Wt[O.N_RED % "w_reduce_c"] + Wt[O.N_RED % "bias_reduce_c"])
        h0 = torch.relu(old_h @ Wt[O.N_RED % "w_reduce_h"] + Wt[O.N_RED % "bias_reduce_h"])
        return out.numpy().astype(self.npdt), (c0.numpy().astype(self.npdt), h0.numpy().astype(self.npdt))

    def encoder_features(self, W, enc_states):
        """attention_decoder.py:66-67."""
        es = self.t(enc_states)
        A = es.shape[-1]
        return (es.reshape(-1, A) @ self.Wt[O.N_WH].reshape(A, A)).reshape(es.shape).numpy()

    def _attention(self, feat, es, mask, c, h, cov, use_cov):
        """attention_decoder.py:79-129 on tensors."""
        Wt = self.Wt
        dec_feat = torch.cat([c, h], dim=1) @ Wt[O.N_ATT_M] + Wt[O.N_ATT_B]
        x = feat + dec_feat[:, None, :]
        if use_cov and cov is not None:
            x = x + cov[:, :, None] * Wt[O.N_WC].reshape(1, 1, -1)
        e = (Wt[O.N_V] * torch.tanh(x)).sum(dim=2)
        a = torch.softmax(e, dim=1) * mask
        a = a / a.sum(dim=1, keepdim=True)
        if use_cov and cov is not None:
            ncov = cov + a
        elif use_cov:
            ncov = a.clone()
        else:
            ncov = None
        ctx = (a[:, :, None] * es).sum(dim=1)
        return ctx, a, ncov

    def _project(self, hps, out, p_gen, attn, ext_ids, n_oov):
        """model.py:228-238, 146-183, 280-285: vocabulary softmax, pointer mixture, top-2K + log."""
        Wt = self.Wt
        vocab_dist = torch.softmax(out @ Wt[O.N_PROJ_W] + Wt[O.N_PROJ_V], dim=1)
        if hps["pointer_gen"]:
            R, V = vocab_dist.shape
            fin = torch.zeros((R, V + n_oov), dtype=self.dt)
            fin[:, :V] = p_gen * vocab_dist
            copy = (1.0 - p_gen) * attn
            ext = torch.as_tensor(np.ascontiguousarray(ext_ids), dtype=torch.int64)
            for r in range(R):                                    # tf.scatter_nd: duplicates accumulate
                fin[r].index_add_(0, ext[r], copy[r])
        else:
            fin = vocab_dist
        fin_np = fin.numpy()
        k2 = 2 * int(hps["batch_size"])
        R = fin_np.shape[0]
        ids = np.zeros((R, k2), dtype=np.int32)
        logp = np.zeros((R, k2), dtype=self.npdt)
        for r in range(R):
            p, i = O.top_k(fin_np[r], k2)                         # tf.nn.top_k's tie rule
            ids[r, :len(i)] = i
            with np.errstate(divide="ignore"):
                logp[r, :len(i)] = np.log(p)
        return ids, logp

    def _tail(self, hps, es, feat, mask, ext_ids, n_oov, tokens, c, h, ctx, cov, use_cov):
        """attention_decoder.py:150-173 after the first attention pass."""
        Wt = self.Wt
        inp = Wt[O.N_EMB][torch.as_tensor(np.asarray(tokens, dtype=np.int64))]
        x = torch.cat([inp, ctx], dim=1) @ Wt[O.N_X_M] + Wt[O.N_X_B]
        c2, h2 = self.lstm_cell(x, c, h, Wt[O.N_DEC_K], Wt[O.N_DEC_B])
        ctx2, attn, _ = self._attention(feat, es, mask, c2, h2, cov, use_cov)
        p_gen = torch.sigmoid(torch.cat([ctx2, c2, h2, x], dim=1) @ Wt[O.N_PG_M] + Wt[O.N_PG_B])
        out = torch.cat([h2, ctx2], dim=1) @ Wt[O.N_OUT_M] + Wt[O.N_OUT_B]
        ids, logp = self._project(hps, out, p_gen, attn, ext_ids, n_oov)
        n = lambda v: None if v is None else v.numpy()
        return ids, logp, n(c2), n(h2), n(attn), n(p_gen[:, 0]), n(cov), n(ctx2)

    def decode_onestep(self, W, hps, enc_states, enc_feat, mask, ext_ids, n_oov, tokens, c, h, prev_cov):
        """model.py:367-443: both attention passes; enc_feat None => W_h.h_i recomputed (what the reference does)."""
        use_cov = bool(hps["coverage"])
        es = self.t(enc_states)
        if enc_feat is None:
            A = es.shape[-1]
            feat = (es.reshape(-1, A) @ self.Wt[O.N_WH].reshape(A, A)).reshape(es.shape)
        else:
            feat = self.t(enc_feat)
        mk, ct, ht = self.t(mask), self.t(c), self.t(h)
        pc = self.t(prev_cov) if (use_cov and prev_cov is not None) else None
        ctx, _, cov = self._attention(feat, es, mk, ct, ht, pc, use_cov)
        return self._tail(hps, es, feat, mk, ext_ids, n_oov, tokens, ct, ht, ctx, cov, use_cov)

    def decode_onestep_dedup(self, W, hps, enc_states, feat, mask, ext_r, n_oov, latest, c, h, cov, hyps, steps):
        """First pass replaced by the parent's carried (context, attention) for steps >= 1."""
        use_cov = bool(hps["coverage"])
        if steps == 0:
            return self.decode_onestep(W, hps, enc_states, feat, mask, ext_r, n_oov, latest, c, h, cov)
        ctx = self.t(np.stack([hh.ctx[0] for hh in hyps]))
        a_prev = self.t(np.stack([hh.ctx[1] for hh in hyps]))
        ncov = self.t(cov) + a_prev if use_cov else None
        return self._tail(hps, self.t(enc_states), self.t(feat), self.t(mask), ext_r, n_oov, latest, self.t(c), self.t(h),
                          ctx, ncov, use_cov)


def decode_ids(W, hps, enc_input, enc_ext, n_oov, mode="literal", dtype=torch.float32, trace: Optional[list] = None):
    """pgen_oracle.decode_ids with the graph side in torch: (tokens after [START], log-probs, avg_log_prob)."""
    if len(enc_input) == 0:
        return [], [], 0.0
    g = TorchGraph(W, dtype)
    saved = (O.run_encoder, O.encoder_features, O.decode_onestep, O._decode_onestep_dedup, O.F32)
    try:
        O.run_encoder, O.encoder_features = g.run_encoder, g.encoder_features
        O.decode_onestep, O._decode_onestep_dedup = g.decode_onestep, g.decode_onestep_dedup
        O.F32 = np.float64 if dtype == torch.float64 else np.float32     # dtype of the host-side state copies
        with torch.no_grad():
            best = O.run_beam_search(W, hps, enc_input, enc_ext, n_oov, mode=mode, trace=trace)
    finally:
        O.run_encoder, O.encoder_features, O.decode_onestep, O._decode_onestep_dedup, O.F32 = saved
    return [int(t) for t in best.tokens[1:]], [float(x) for x in best.log_probs[1:]], best.avg_log_prob
