// Launchers of the sm_100a kernels of the decode path.  Every launcher enqueues on `st` and
// returns the number of kernels it launched (for pgen_stats.kernel_launches).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <mutex>

namespace pgn {

// Function attributes (dynamic shared memory limits) are per device and per process: handles on several GPUs, driven
// from several threads of one process (Flink subtasks of one TaskManager), set them exactly once each.
struct PerDeviceOnce {
  std::mutex mu;
  size_t done[64] = {};
  template <class F> void ensure(size_t need, F&& set) {
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(mu);
    if (done[dev & 63] < need) { set(); done[dev & 63] = need; }
  }
};

// C[M,N] = act( [A0 | A1][M, k0+k1] . B[N, k0+k1]^T + bias[N] ), fp32 SIMT tiles.
// A0 rows may be gathered through gather0 (row m reads A0 + gather0[m]*lda0).
struct GemmArgs {
  const float* A0 = nullptr; int lda0 = 0; int k0 = 0; const int* gather0 = nullptr;
  const float* A1 = nullptr; int lda1 = 0; int k1 = 0;
  const float* B = nullptr;  int ldb = 0;
  const float* bias = nullptr;
  float* C = nullptr; int ldc = 0;
  float* C_tf32 = nullptr; int ldc_tf32 = 0;   // optional second output: C rounded to tf32 (rna), operand of the tensor-core pass
  float* C_pack = nullptr; int pack_K = 0;     // optional second output (N == 512): 2 log2(e) * C in the layout attn_fused.cu
                                               // copies, [row / K][(row % K) / 2][n][row % 2] (beam rows interleaved in pairs)
  int M = 0, N = 0;
  int act = 0;            // 0 none, 1 relu
  int skinny = 0;         // per-step decoder GEMM (few hundred rows): split-K tile kernel.  Chosen by the call
                          // site, never by M, so that an article's rounding does not depend on the batch size
};
int launch_gemm_nt(const GemmArgs& a, cudaStream_t st);

struct Dims {
  int H, E, A, V;         // hidden, emb, 2H, vocab
  int K, K2;              // beam, 2*beam
  int T, S;               // max_enc_steps (row stride of [.,T] buffers), max_dec_steps
  int SW;                 // state row width A + 2H + E : [ctx | c | h | x]
  int min_dec_steps;
  int use_cov, pointer_gen;
};

// ---- encoder ----
// One time step of both LSTM directions for every article (model.py:76-94).  zx [2][B*T][4H]
// holds x.Kx + bias; kh [2][H][4H]; h state double-buffered by the caller.
int launch_enc_lstm_step(const Dims& d, int B, int s, const int* lens, const float* zx, const float* kh,
                         const float* h_prev, float* h_next, float* c_state, float* enc_states,
                         cudaStream_t st);

// The whole recurrence in one launch (hidden_dim == 256 only): clusters of 8 CTAs, recurrent weights resident in
// shared memory, h exchanged through DSMEM.  Writes the final (h, c) of both directions to h_fin / c_fin [2][B][H].
// ids == nullptr: zx is the batch's input projection [2][B][T][4H];  ids != nullptr: zx is the per-word table
// [2][V][4H] and ids [B][T] are the encoder tokens.
int launch_enc_lstm_persistent(const Dims& d, int B, const int* lens, const float* zx, const int* ids, const float* kh,
                               float* h_fin, float* c_fin, float* enc_states, cudaStream_t st);

// ---- decoder step ----
struct StepBufs {
  const int* lens; const int* n_oov;
  const int* active;                    // [B]
  const int* sel_parent; const int* sel_token;   // [B,K]
  const float* emb;                     // [V,E]
  const float* state_old; float* state_new;      // [R,SW]
  const float* attn_old; float* attn_new;        // [R,T]
  const float* cov_old; float* cov_new;          // [R,T]
  float* xin;                           // [R,E+A]
  float* hprev; float* cprev;           // [R,H]
};
int launch_gather_step(const Dims& d, int B, const StepBufs& b, cudaStream_t st);
int launch_lstm_pointwise(const Dims& d, int R, const float* z, const float* cprev, float* state_new,
                          const int* active, cudaStream_t st);
// energies e[r,t] = sum_a v[a] tanh(F[b,t,a] + dfeat[r,a] + cov[r,t] w_c[a])   (attention_decoder.py:105-117)
// rows_per_article: K in the step loop, 1 for the initial pass (row b*K only).
int launch_attn_energy(const Dims& d, int B, int rows_per_article, const int* lens, const int* active,
                       const float* enc_feat, const float* dfeat, const float* cov, const float* v,
                       const float* w_c, float* e, cudaStream_t st);
// softmax, mask, renormalise, context (attention_decoder.py:96-101, 126-127); writes attn [R,T] and
// ctx into state[r, 0:A].
int launch_attn_context(const Dims& d, int B, int rows_per_article, const int* lens, const int* active,
                        const float* enc_states, const float* e, float* attn, float* state, cudaStream_t st);
// The same attention pass (all K rows of every live article) in ONE launch, hidden_dim 256 and beam 4 or 8 only
// (attn_fused.cu): persistent CTAs stream the encoder features and states through shared memory once and leave
// per-segment partials in `pctx` / `segstat`, which a second small kernel merges per article (`ticket`: one int, zero
// before the launch and left zero by it).  ctx of row r goes to ctx + r * ctx_ld.
bool attn_fused_supported(const Dims& d);
int attn_fused_max_articles();                       // per launch (articles_per_batch must not exceed it)
int attn_fused_nseg(const Dims& d);
size_t attn_fused_pctx_floats(const Dims& d, int B);
size_t attn_fused_segstat_floats(const Dims& d, int B);
cudaError_t attn_fused_prepare(const Dims& d);       // per device, before the first launch (and before any capture)
// dpack [B][K/2][512][2] = 2 log2(e) * dfeat with the article's beam rows interleaved in pairs: what the kernel's
// producer copies into shared memory in one piece (the decoder cell writes this layout directly)
int launch_attn_pack_dfeat(const Dims& d, int B, const float* dfeat, float* dpack, const int* active, cudaStream_t st);
int launch_attn_fused(const Dims& d, int B, const int* lens, const int* active, const float* enc_feat,
                      const float* enc_states, const float* dpack, const float* cov, const float* v, const float* w_c,
                      float* e, float* pctx, float* segstat, int* ticket, float* attn, float* ctx, int ctx_ld, int n_sms,
                      cudaStream_t st);
// p_gen = sigmoid(state[r,:] . pg_w + pg_b)                                     (attention_decoder.py:166-167)
int launch_pgen(const Dims& d, int R, const float* state, const float* pg_w, const float* pg_b, float* pgen,
                const int* active, cudaStream_t st);
// per-row max and sum(exp(l - max)) over the V logits                          (model.py:238)
int launch_softmax_stats(const Dims& d, int R, const float* logits, float* rowmax, float* rowsum,
                         const int* active, cudaStream_t st);
// final distribution (model.py:146-183) restricted to what top_k can return, top-2K + log (model.py:284-285)
struct FinalArgs {
  const float* logits; const float* rowmax; const float* rowsum; const float* pgen; const float* attn;
  const int* uniq_ids; const int* uniq_cnt; const int* pos_start; const int* pos_list;   // per article
  const int* n_oov; const int* active;
  int* topk_ids; float* topk_logp;      // [R,2K]
};
int launch_final_topk(const Dims& d, int B, const FinalArgs& a, cudaStream_t st);

struct BeamBufs {
  int* active; int* sel_parent; int* sel_token; double* beam_sum;       // [B],[B,K],[B,K],[B,K]
  int* res_count; double* res_sum; int* res_len; int* res_step; int* res_parent; float* res_lp;   // [B],[B,K]...
  int* hist_tok; int* hist_par; float* hist_logp;                       // [B,S,K]
  int* steps_done; int* n_active;                                       // [B], [1]
  int* best_kind; int* best_idx;                                        // [B] 0: result idx, 1: live beam slot
};
// run_beam_search's per-step bookkeeping (beam_search.py:126-156), one block per article.
int launch_beam_step(const Dims& d, int B, const int* topk_ids, const float* topk_logp,
                     const BeamBufs& b, cudaStream_t st);
// beam_search.py:160-168 + token back-trace of the winning hypothesis.
int launch_finalize(const Dims& d, int B, const int* lens, const BeamBufs& b, int* out_tokens, int* out_len,
                    float* out_logp, double* out_avg, cudaStream_t st);
int launch_init_beam(const Dims& d, int B, const int* lens, const BeamBufs& b, cudaStream_t st);
// replicate the initial decoder state of row b*K into the "old" buffers (beam_search.py:100-106)
int launch_fill(float* p, float v, size_t n, cudaStream_t st);

}  // namespace pgn
