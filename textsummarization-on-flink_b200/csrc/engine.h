// Device-side decode engine: weights in HBM, one batch slot, encoder + beam-search step loop.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "kernels.h"
#include "text.h"
#include "vocab_tc.h"
#include "dec_cell.h"

namespace pgn {

struct DeviceWeights {
  float *emb, *enc_kxT[2], *enc_kh[2], *enc_b[2];
  float *red_wcT, *red_whT, *red_bc, *red_bh;
  float *WhT, *v, *w_c, *att_MT, *att_b, *x_MT, *x_b, *dec_kT, *dec_b, *pg_w, *pg_b, *out_MT, *out_b;
  float *proj_wT, *proj_v;
  float proj_wnorm_max = 0.f;      // max over the vocabulary of the L2 norm of a word's projection column
};

struct EngineStats {
  uint64_t batches = 0, decode_steps = 0, article_steps = 0, early_stops = 0, kernel_launches = 0;
  double ms_h2d = 0, ms_encoder = 0, ms_decoder = 0, ms_d2h = 0;
  double ms_enc_recurrence = 0, ms_enc_features = 0;    // parts of ms_encoder: BiLSTM recurrence, W_h.h_i product
  double ms_attention = 0, ms_vocab = 0, ms_beam = 0;   // summed over the steps of the last profiled batch
  uint64_t prof_steps = 0;
  uint64_t selection_widened = 0, selection_fallbacks = 0;   // rows whose tf32 candidate set failed the exactness guard
};

class Engine {
 public:
  ~Engine();
  // Returns "" or an error message; *code receives the PGEN_ERR_* class.
  // `share` (optional): another engine of the same handle whose weight replica (fp32 + tf32 copy) is reused, so
  // that all slots of a handle read ONE copy of the weights.
  std::string init(const HyperParams& hp, int vocab_size, int device, int* code, const Engine* share = nullptr);
  std::string upload(int n, const int32_t* enc_ids, const int32_t* ext_ids, int ld, const int32_t* lens,
                     const int32_t* n_oovs);
  std::string run();
  std::string sync();
  bool done();                       // the last run() (including its D2H copies) has finished
  // expect_rows >= 0: refuse unless that many articles were staged (the caller sized its buffers for them)
  std::string download(int expect_rows, int32_t* out_tokens, int32_t* out_lens, float* out_logp, double* out_avg);
  std::string fetch(const std::string& name, void* out, size_t cap_bytes, size_t* n);
  std::string set_option(const std::string& kv);
  std::string vocab_logits_op(int rows, const float* out_dev, float* logits_dev, int use_tc);
  std::string attention_op(int n, const float* enc_feat, const float* enc_states, const int32_t* lens,
                           const float* dfeat, const float* cov, float* attn, float* ctx);
  const DeviceWeights& weights() const { return w_; }
  int device() const { return device_; }
  cudaStream_t stream() const { return st_; }
  int capacity() const { return Bcap_; }
  const Dims& dims() const { return d_; }
  EngineStats stats;

 private:
  std::string load_weights(const std::string& path);
  void encoder();
  void decoder();
  template <class T> T* dalloc(size_t n);
  std::string check(const char* what);

  Dims d_{};
  int device_ = 0, Bcap_ = 0, R_ = 0, nB_ = 0, max_len_ = 0;
 public:
  volatile int where_ = 0;                 // host-side breadcrumb (pgen_debug_where): which call the engine's thread is inside
  void debug_progress(long long* out3);
  void debug_live(long long* out48);       // kernel entry / exit counters of this engine's device (PGEN_LIVE builds)    // (live articles, steps done by article 0, attention ticket) read on a stream of its own
 private:
  bool staged_ = false, ran_ = false, synced_ = false, counted_ = false;
  bool use_graph_ = true;            // replay captured steps instead of launching ~13 kernels per step
  int graph_steps_ = 10;             // steps per graph (even); the early-stop poll happens between graphs
  int graph_chunk_ = 0, graph_nodes_ = 0;
  bool graph_failed_ = false;
  std::map<int, cudaGraphExec_t> graphs_;     // by articles in the batch
  std::string measure(int reps);
  int step_launch(int B, int o, int ps);
  cudaGraphExec_t step_graph(int B, int chunk, bool capture);
  int check_every_ = 0;              // > 0: poll the device every N steps (after min_dec_steps) to stop early;
                                     // 0: never block the host -- run() is fully asynchronous
  bool profile_ = false;
  int profile_first_ = 0;            // profile=N (N > 1): per-step sums over the first N steps only
  int vocab_guard_ = 0;              // MergeArgs::guard
  int* sel_stats_ = nullptr;         // device [2], cumulative
  int32_t* h_selstats_ = nullptr;    // pinned copy, refreshed with every batch's results
  bool enc_persistent_ = true;
  DecCell dcell_;
  bool dec_fused_ = false;           // one-launch decoder cell + out projection / p_gen (dec_cell.cu) instead of gather + GEMMs
  bool dpack_fresh_ = false;         // the decoder cell has written dpack_ for the coming attention pass
  bool attn_fused_ = false;          // one-launch attention (attn_fused.cu) instead of the energy + context pair
  unsigned char* flush_ = nullptr;   // 256 MB written between the replays of measure() so that they start from a cold L2
  float *att_pctx_ = nullptr, *att_segstat_ = nullptr, *dpack_ = nullptr;
  int* att_cnt_ = nullptr;
  int attention(int B, const float* cov, float* attn, float* state);   // the step's attention pass on dfeat_
  bool vocab_tc_ = false;          // tensor-core vocabulary path in use (falls back to the SIMT kernels when
                                   // the dims are outside what vocab_tc.cu covers)
  VocabTC vtc_;
  std::string vtc_why_;
  float *proj_wT_tf32_ = nullptr, *out_tf32_ = nullptr, *part_m_ = nullptr, *part_s_ = nullptr, *artl_ = nullptr;
  float4* gmax_ = nullptr;
  int gstride_ = 0;
  float* smax_ = nullptr;
  int sstride_ = 0;
  int n_sms_ = 148;
  std::vector<cudaEvent_t> pev_;   // 4 events per step when profiling: attn begin, attn end/vocab begin... 
  int pev_used_ = 0;
  cudaStream_t st_ = nullptr;
  cudaStream_t st2_ = nullptr;       // side branch of a decode step (p_gen, article-word logits)
  cudaEvent_t fev_[3] = {nullptr, nullptr, nullptr};
  cudaEvent_t ev_[7] = {};
  cudaEvent_t eev_[4] = {};          // encoder: recurrence begin / end, features begin / end
  DeviceWeights w_{};
  std::vector<void*> allocs_;
  // host staging (pinned)
  int32_t *h_ids_ = nullptr, *h_ext_ = nullptr, *h_lens_ = nullptr, *h_noov_ = nullptr, *h_uniq_ = nullptr,
          *h_ucnt_ = nullptr, *h_pstart_ = nullptr, *h_plist_ = nullptr, *h_nactive_ = nullptr;
  int32_t *h_out_tok_ = nullptr, *h_out_len_ = nullptr, *h_steps_ = nullptr, *h_rescnt_ = nullptr;
  float* h_out_lp_ = nullptr;
  double* h_out_avg_ = nullptr;
  // device batch slot
  int *ids_, *ext_, *lens_, *noov_, *uniq_, *ucnt_, *pstart_, *plist_;
  float *wh_hi_ = nullptr, *wh_lo_ = nullptr, *es_hi_ = nullptr, *es_lo_ = nullptr;   // 3xTF32 operands of W_h . h_i
  GemmTC3 g3_;
  bool feat_tc_ = false;
  float* zx_tab_ = nullptr;          // [2][V][4H] input projection of every word (shared by the handle's engines)
  float *zx_, *enc_states_, *enc_feat_, *h_enc_[2], *c_enc_;
  float *state_[2], *attn_[2], *cov_[2], *xin_, *hprev_, *cprev_, *z_, *dfeat_, *e_, *out_, *pgen_, *logits_,
      *rowmax_, *rowsum_, *topk_logp_;
  int* topk_ids_;
  BeamBufs bb_{};
  int *out_tok_, *out_len_;
  float* out_lp_;
  double* out_avg_;
};

}  // namespace pgn
