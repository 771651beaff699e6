#include "engine.h"

#include <algorithm>
#include <cstdio>
#include <cmath>
#include <cstring>
#include <functional>
#include <cstdlib>
#include <fstream>

#include "../../include/pgen.h"

namespace pgn {

#define CK(call)                                                                        \
  do {                                                                                  \
    cudaError_t e__ = (call);                                                           \
    if (e__ != cudaSuccess) return std::string(#call) + ": " + cudaGetErrorString(e__); \
  } while (0)

std::string Engine::check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return std::string(what) + ": " + cudaGetErrorString(e);
  return "";
}

template <class T>
T* Engine::dalloc(size_t n) {
  void* p = nullptr;
  if (cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)) != cudaSuccess) return nullptr;
  cudaMemset(p, 0, std::max<size_t>(n, 1) * sizeof(T));
  allocs_.push_back(p);
  return (T*)p;
}

Engine::~Engine() {
  if (device_ >= 0) cudaSetDevice(device_);
  for (auto& kv : graphs_) cudaGraphExecDestroy(kv.second);
  for (void* p : allocs_) cudaFree(p);
  for (void* p : {(void*)h_ids_, (void*)h_ext_, (void*)h_lens_, (void*)h_noov_, (void*)h_uniq_, (void*)h_ucnt_,
                  (void*)h_pstart_, (void*)h_plist_, (void*)h_nactive_, (void*)h_out_tok_, (void*)h_out_len_,
                  (void*)h_out_lp_, (void*)h_out_avg_, (void*)h_steps_, (void*)h_rescnt_, (void*)h_selstats_})
    if (p) cudaFreeHost(p);
  for (auto& e : ev_) if (e) cudaEventDestroy(e);
  for (auto& e : eev_) if (e) cudaEventDestroy(e);
  for (auto& e : pev_) if (e) cudaEventDestroy(e);
  if (st2_) cudaStreamDestroy(st2_);
  for (auto& e : fev_) if (e) cudaEventDestroy(e);
  if (st_) cudaStreamDestroy(st_);
}

// --------------------------------------------------------------------------------------
// .pgenw archive (see weights_io.py) -> device, every matrix stored [N_out, K_in]
// --------------------------------------------------------------------------------------
namespace {
struct HostTensor { std::vector<uint32_t> dims; const float* data; size_t n; };

std::string read_archive(const std::string& path, std::vector<char>* buf, std::map<std::string, HostTensor>* out) {
  std::ifstream f(path, std::ios::binary | std::ios::ate);
  if (!f) return "cannot open weights archive '" + path + "'";
  const size_t sz = (size_t)f.tellg();
  buf->resize(sz);
  f.seekg(0);
  f.read(buf->data(), sz);
  if (sz < 12 || memcmp(buf->data(), "PGENW001", 8) != 0) return "'" + path + "' is not a .pgenw archive";
  uint32_t n;
  memcpy(&n, buf->data() + 8, 4);
  size_t p = 12;
  for (uint32_t i = 0; i < n; ++i) {
    if (p + 2 > sz) return "truncated weights archive";
    uint16_t ln;
    memcpy(&ln, buf->data() + p, 2); p += 2;
    if (p + ln + 2 > sz) return "truncated weights archive";
    std::string name(buf->data() + p, ln); p += ln;
    uint8_t dtype = (uint8_t)(*buf)[p], ndim = (uint8_t)(*buf)[p + 1]; p += 2;
    if (dtype != 0) return "tensor '" + name + "': unsupported dtype";
    HostTensor t;
    t.dims.resize(ndim);
    if (p + 4 * ndim + 16 > sz) return "truncated weights archive";
    if (ndim) memcpy(t.dims.data(), buf->data() + p, 4 * ndim);
    p += 4 * ndim;
    uint64_t off, nb;
    memcpy(&off, buf->data() + p, 8); memcpy(&nb, buf->data() + p + 8, 8); p += 16;
    if (off + nb > sz || off % 4) return "tensor '" + name + "': bad offset";
    t.data = reinterpret_cast<const float*>(buf->data() + off);
    t.n = nb / 4;
    (*out)[name] = t;
  }
  return "";
}
}  // namespace

std::string Engine::load_weights(const std::string& path) {
  std::vector<char> buf;
  std::map<std::string, HostTensor> ts;
  std::string err = read_archive(path, &buf, &ts);
  if (!err.empty()) return err;
  const int H = d_.H, E = d_.E, A = d_.A, V = d_.V;
  const std::string P = "seq2seq/", D = P + "decoder/attention_decoder/";
  auto need = [&](const std::string& name, size_t n, const float** out) -> std::string {
    auto it = ts.find(name);
    if (it == ts.end()) return "weights archive lacks variable '" + name + "'";
    if (it->second.n != n) {
      char m[256];
      snprintf(m, sizeof m, "variable '%s' has %zu elements, the flags imply %zu", name.c_str(), it->second.n, n);
      return m;
    }
    *out = it->second.data;
    return "";
  };
  // upload helpers: as-is, or transposed [rows, cols] -> [cols, rows], optionally a row range first
  auto up = [&](const float* src, size_t n, float** dst) -> std::string {
    *dst = dalloc<float>(n);
    if (!*dst) return "cudaMalloc failed for weights";
    CK(cudaMemcpy(*dst, src, n * sizeof(float), cudaMemcpyHostToDevice));
    return "";
  };
  auto upT = [&](const float* src, int rows, int cols, float** dst) -> std::string {
    std::vector<float> t((size_t)rows * cols);
    for (int r = 0; r < rows; ++r)
      for (int c = 0; c < cols; ++c) t[(size_t)c * rows + r] = src[(size_t)r * cols + c];
    return up(t.data(), t.size(), dst);
  };
  const float* p;
#define NEED(name, n) do { err = need(name, n, &p); if (!err.empty()) return err; } while (0)
#define UP(dst, n) do { err = up(p, n, dst); if (!err.empty()) return err; } while (0)
#define UPT(dst, r, c) do { err = upT(p, r, c, dst); if (!err.empty()) return err; } while (0)
  NEED(P + "embedding/embedding", (size_t)V * E); UP(&w_.emb, (size_t)V * E);
  const char* dn[2] = {"fw", "bw"};
  for (int dir = 0; dir < 2; ++dir) {
    const std::string b = P + "encoder/bidirectional_rnn/" + dn[dir] + "/lstm_cell/";
    NEED(b + "kernel", (size_t)(E + H) * 4 * H);
    const float* k = p;
    p = k; UPT(&w_.enc_kxT[dir], E, 4 * H);                       // rows 0..E-1 : x part  -> [4H, E]
    p = k + (size_t)E * 4 * H; UP(&w_.enc_kh[dir], (size_t)H * 4 * H);   // rows E.. : h part, kept [H, 4H]
    NEED(b + "bias", (size_t)4 * H); UP(&w_.enc_b[dir], (size_t)4 * H);
  }
  // the two directions' recurrent kernels must be contiguous for the step kernel: re-pack
  {
    float* both = dalloc<float>((size_t)2 * H * 4 * H);
    if (!both) return "cudaMalloc failed for weights";
    for (int dir = 0; dir < 2; ++dir)
      CK(cudaMemcpy(both + (size_t)dir * H * 4 * H, w_.enc_kh[dir], (size_t)H * 4 * H * sizeof(float), cudaMemcpyDeviceToDevice));
    w_.enc_kh[0] = both; w_.enc_kh[1] = both + (size_t)H * 4 * H;
  }
  NEED(P + "reduce_final_st/w_reduce_c", (size_t)A * H); UPT(&w_.red_wcT, A, H);
  NEED(P + "reduce_final_st/w_reduce_h", (size_t)A * H); UPT(&w_.red_whT, A, H);
  NEED(P + "reduce_final_st/bias_reduce_c", (size_t)H); UP(&w_.red_bc, (size_t)H);
  NEED(P + "reduce_final_st/bias_reduce_h", (size_t)H); UP(&w_.red_bh, (size_t)H);
  NEED(D + "W_h", (size_t)A * A); UPT(&w_.WhT, A, A);
  NEED(D + "v", (size_t)A); UP(&w_.v, (size_t)A);
  if (d_.use_cov) { NEED(D + "coverage/w_c", (size_t)A); UP(&w_.w_c, (size_t)A); }
  else w_.w_c = w_.v;
  NEED(D + "Attention/Linear/Matrix", (size_t)2 * H * A); UPT(&w_.att_MT, 2 * H, A);
  NEED(D + "Attention/Linear/Bias", (size_t)A); UP(&w_.att_b, (size_t)A);
  NEED(D + "Linear/Matrix", (size_t)(E + A) * E); UPT(&w_.x_MT, E + A, E);
  NEED(D + "Linear/Bias", (size_t)E); UP(&w_.x_b, (size_t)E);
  NEED(D + "lstm_cell/kernel", (size_t)(E + H) * 4 * H); UPT(&w_.dec_kT, E + H, 4 * H);
  NEED(D + "lstm_cell/bias", (size_t)4 * H); UP(&w_.dec_b, (size_t)4 * H);
  if (d_.pointer_gen) {
    NEED(D + "calculate_pgen/Linear/Matrix", (size_t)d_.SW); UP(&w_.pg_w, (size_t)d_.SW);
    NEED(D + "calculate_pgen/Linear/Bias", (size_t)1); UP(&w_.pg_b, (size_t)1);
  }
  NEED(D + "AttnOutputProjection/Linear/Matrix", (size_t)(H + A) * H); UPT(&w_.out_MT, H + A, H);
  NEED(D + "AttnOutputProjection/Linear/Bias", (size_t)H); UP(&w_.out_b, (size_t)H);
  NEED(P + "output_projection/w", (size_t)H * V); UPT(&w_.proj_wT, H, V);
  {                                                  // largest column norm: error bound of the tf32 logits (vocab_tc.cu)
    std::vector<double> sq((size_t)V, 0.0);
    for (int h = 0; h < H; ++h)
      for (int v = 0; v < V; ++v) { const double x = p[(size_t)h * V + v]; sq[v] += x * x; }
    double mx = 0.0;
    for (int v = 0; v < V; ++v) mx = std::max(mx, sq[v]);
    w_.proj_wnorm_max = (float)(std::sqrt(mx) * (1.0 + 1e-6));
  }
  NEED(P + "output_projection/v", (size_t)V); UP(&w_.proj_v, (size_t)V);
#undef NEED
#undef UP
#undef UPT
  return "";
}

std::string Engine::init(const HyperParams& hp, int vocab_size, int device, int* code, const Engine* share) {
  *code = PGEN_ERR_CUDA;
  device_ = device;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return "no CUDA device available (libpgen_cuda has no CPU fallback)";
  }
  if (device < 0 || device >= ndev) {
    *code = PGEN_ERR_ARG;
    return "device ordinal " + std::to_string(device) + " out of range (" + std::to_string(ndev) + " visible)";
  }
  CK(cudaSetDevice(device));
  *code = PGEN_ERR_UNSUPPORTED;
  d_.H = (int)hp.hidden_dim; d_.E = (int)hp.emb_dim; d_.A = 2 * d_.H; d_.V = vocab_size;
  d_.K = (int)hp.beam_size; d_.K2 = 2 * d_.K; d_.T = (int)hp.max_enc_steps; d_.S = (int)hp.max_dec_steps;
  d_.SW = d_.A + 2 * d_.H + d_.E; d_.min_dec_steps = (int)hp.min_dec_steps;
  d_.use_cov = hp.coverage ? 1 : 0; d_.pointer_gen = hp.pointer_gen ? 1 : 0;
  if (d_.H < 32 || d_.H % 32) return "hidden_dim must be a positive multiple of 32";
  if (d_.E < 16 || d_.E % 16) return "emb_dim must be a positive multiple of 16";
  if (d_.K < 1 || d_.K > 16) return "beam_size must be in 1..16";
  if (d_.T < 1 || d_.S < 1) return "max_enc_steps and max_dec_steps must be positive";
  if (d_.V < 2 * d_.K + 4) return "vocabulary smaller than 2*beam_size";
  if (hp.articles_per_batch < 1 || hp.articles_per_batch > 4096) { *code = PGEN_ERR_ARG; return "articles_per_batch must be in 1..4096"; }
  Bcap_ = (int)hp.articles_per_batch;
  R_ = Bcap_ * d_.K;
  *code = PGEN_ERR_CUDA;
  CK(cudaStreamCreateWithFlags(&st_, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&st2_, cudaStreamNonBlocking));
  for (auto& e : fev_) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  for (auto& e : ev_) CK(cudaEventCreate(&e));
  for (auto& e : eev_) CK(cudaEventCreate(&e));

  *code = PGEN_ERR_IO;
  std::string wpath = hp.weights_path;
  if (wpath.empty()) wpath = hp.log_root + "/" + hp.exp_name + "/train/model.pgenw";   // util.py:34
  std::string err;
  if (share) w_ = share->w_;                       // borrowed device pointers; the owner frees them
  else err = load_weights(wpath);
  if (!err.empty()) {
    if (err.find("cudaM") != std::string::npos) *code = PGEN_ERR_CUDA;   // a failed cudaMalloc / cudaMemcpy is not an I/O error
    return err;
  }

  *code = PGEN_ERR_CUDA;
  const size_t B = Bcap_, T = d_.T, K = d_.K, R = R_, H = d_.H, A = d_.A, E = d_.E, S = d_.S, V = d_.V;
#define HOSTALLOC(p, n) CK(cudaMallocHost((void**)&p, std::max<size_t>((n), 1) * sizeof(*p)))
  HOSTALLOC(h_ids_, B * T); HOSTALLOC(h_ext_, B * T); HOSTALLOC(h_lens_, B); HOSTALLOC(h_noov_, B);
  HOSTALLOC(h_uniq_, B * T); HOSTALLOC(h_ucnt_, B); HOSTALLOC(h_pstart_, B * (T + 1)); HOSTALLOC(h_plist_, B * T);
  HOSTALLOC(h_nactive_, 1); HOSTALLOC(h_out_tok_, B * S); HOSTALLOC(h_out_len_, B); HOSTALLOC(h_out_lp_, B * S);
  HOSTALLOC(h_out_avg_, B); HOSTALLOC(h_steps_, B); HOSTALLOC(h_rescnt_, B); HOSTALLOC(h_selstats_, 2);
  h_selstats_[0] = h_selstats_[1] = 0;
#undef HOSTALLOC
#define DALLOC(p, type, n) do { p = dalloc<type>(n); if (!p) return std::string("cudaMalloc failed for " #p); } while (0)
  DALLOC(ids_, int, B * T); DALLOC(ext_, int, B * T); DALLOC(lens_, int, B); DALLOC(noov_, int, B);
  DALLOC(uniq_, int, B * T); DALLOC(ucnt_, int, B); DALLOC(pstart_, int, B * (T + 1)); DALLOC(plist_, int, B * T);
  DALLOC(zx_, float, 2 * B * T * 4 * H); DALLOC(enc_states_, float, B * T * A); DALLOC(enc_feat_, float, B * T * A);
  DALLOC(h_enc_[0], float, 2 * B * H); DALLOC(h_enc_[1], float, 2 * B * H); DALLOC(c_enc_, float, 2 * B * H);
  for (int i = 0; i < 2; ++i) {
    DALLOC(state_[i], float, R * d_.SW); DALLOC(attn_[i], float, R * T); DALLOC(cov_[i], float, R * T);
  }
  DALLOC(xin_, float, R * (E + A)); DALLOC(hprev_, float, R * H); DALLOC(cprev_, float, R * H);
  DALLOC(z_, float, R * 4 * H); DALLOC(dfeat_, float, R * A); DALLOC(e_, float, R * T); DALLOC(out_, float, R * H);
  DALLOC(pgen_, float, R); DALLOC(logits_, float, R * V); DALLOC(rowmax_, float, R); DALLOC(rowsum_, float, R);
  DALLOC(topk_ids_, int, R * 2 * K); DALLOC(topk_logp_, float, R * 2 * K);
  DALLOC(bb_.active, int, B); DALLOC(bb_.sel_parent, int, R); DALLOC(bb_.sel_token, int, R);
  DALLOC(bb_.beam_sum, double, R); DALLOC(bb_.res_count, int, B); DALLOC(bb_.res_sum, double, R);
  DALLOC(bb_.res_len, int, R); DALLOC(bb_.res_step, int, R); DALLOC(bb_.res_parent, int, R); DALLOC(bb_.res_lp, float, R);
  DALLOC(bb_.hist_tok, int, B * S * K); DALLOC(bb_.hist_par, int, B * S * K); DALLOC(bb_.hist_logp, float, B * S * K);
  DALLOC(bb_.steps_done, int, B); DALLOC(bb_.n_active, int, 1); DALLOC(bb_.best_kind, int, B); DALLOC(bb_.best_idx, int, B);
  DALLOC(out_tok_, int, B * S); DALLOC(out_len_, int, B); DALLOC(out_lp_, float, B * S); DALLOC(out_avg_, double, B);
  {
    cudaDeviceGetAttribute(&n_sms_, cudaDevAttrMultiProcessorCount, device_);
    const size_t Rpad = (R + 127) / 128 * 128;
    if (share) proj_wT_tf32_ = share->proj_wT_tf32_;
    else DALLOC(proj_wT_tf32_, float, V * H);
    DALLOC(out_tf32_, float, Rpad * H);
    const size_t n_part = 2 * ((V + 255) / 256);                   // one (max, sum) per row and 128 words
    DALLOC(part_m_, float, n_part * R); DALLOC(part_s_, float, n_part * R);
    gstride_ = (int)((((V + 31) / 32) + 7) & ~(size_t)7);      // float4 per row, rows 128-byte aligned
    DALLOC(gmax_, float4, (size_t)gstride_ * Rpad);
    sstride_ = (int)((((V + 255) / 256) * 2 + 31) & ~(size_t)31);
    DALLOC(smax_, float, (size_t)sstride_ * Rpad);
    DALLOC(artl_, float, R * T);
    DALLOC(sel_stats_, int, 2);
    if (attn_fused_supported(d_) && (int)B <= attn_fused_max_articles()) {
      DALLOC(att_pctx_, float, attn_fused_pctx_floats(d_, (int)B)); DALLOC(att_segstat_, float, attn_fused_segstat_floats(d_, (int)B));
      DALLOC(dpack_, float, R * A);
      DALLOC(att_cnt_, int, B + 1);                                  // per-article segment counters + the ticket counter
      attn_fused_ = attn_fused_prepare(d_) == cudaSuccess;
      if (!attn_fused_) cudaGetLastError();
    }
    if (!share) launch_round_tf32(w_.proj_wT, proj_wT_tf32_, V * H, st_);
    if (H == 256 && hp.encoder_table) {
      // The encoder's input projection depends on the token only: x.Kx + b of every vocabulary word, once per model
      // (same GEMM kernel, so the rows are bit-identical to the per-batch projection), read by the recurrence by id.
      if (share) zx_tab_ = share->zx_tab_;
      else {
        DALLOC(zx_tab_, float, 2 * V * 4 * H);
        for (int dir = 0; dir < 2; ++dir) {
          GemmArgs g;
          g.A0 = w_.emb; g.lda0 = (int)E; g.k0 = (int)E; g.B = w_.enc_kxT[dir]; g.ldb = (int)E; g.bias = w_.enc_b[dir];
          g.C = zx_tab_ + (size_t)dir * V * 4 * H; g.ldc = (int)(4 * H); g.M = (int)V; g.N = (int)(4 * H);
          launch_gemm_nt(g, st_);
        }
      }
    }
    if (A % 256 == 0) {
      // W_h . h_i on the tensor cores with fp32 accuracy (3xTF32): hi/lo parts of the weights once per model, of the
      // encoder states once per batch
      const size_t A2 = A * A, n_es = B * T * A;
      if (share) { wh_hi_ = share->wh_hi_; wh_lo_ = share->wh_lo_; }
      else {
        DALLOC(wh_hi_, float, A2); DALLOC(wh_lo_, float, A2);
        launch_split_tf32(w_.WhT, wh_hi_, wh_lo_, A2, st_);
      }
      DALLOC(es_hi_, float, n_es); DALLOC(es_lo_, float, n_es);
      std::string why = g3_.init(es_hi_, es_lo_, (int)(B * T), wh_hi_, wh_lo_, (int)A, (int)A, n_sms_);
      feat_tc_ = g3_.ready;
      (void)why;
    }
    if (H == 256 && E == 128) {
      if (share) dcell_.share(share->dcell_);
      else dcell_.init(d_, w_.x_MT, w_.dec_kT, w_.att_MT, w_.out_MT, [&](size_t nf) { return dalloc<float>(nf); }, st_);
      // One batch in flight: the one-launch cell is 15 % faster end to end (latency).  Several batches in flight: the GPU
      // is occupied in SM-time, the cell holds 128 SMs exclusively (203 KB of shared memory per CTA) while the split-K
      // GEMMs of one batch share SMs with the kernels of the others, and the split path decodes 6 % more articles/s
      // (tools/slot_sweep.py).  --decoder=fused|split overrides.
      dec_fused_ = dcell_.ready && (hp.decoder == "fused" || (hp.decoder == "auto" && hp.slots == 1));
    }
    if (V >= 128) vtc_why_ = vtc_.init(proj_wT_tf32_, (int)V, (int)H, out_tf32_, (int)Rpad, 2 * (int)K, n_sms_);
    else vtc_why_ = "vocabulary smaller than one 128-word tile";
    vocab_tc_ = vtc_.ready;
  }
#undef DALLOC
  {
    // Dynamic shared memory of the kernels whose footprint grows with max_enc_steps / beam_size / vocab_size: refuse
    // at create, with the numbers, rather than fail at the first batch with a bare launch error.
    int optin = 0;
    CK(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device_));
    const size_t KM = d_.K <= 4 ? 4 : d_.K <= 8 ? 8 : 16, K2M = d_.K2 <= 8 ? 8 : d_.K2 <= 16 ? 16 : 32;
    struct { const char* what; size_t bytes; } need[] = {
        {"attn_context_kernel (softmax + context, max_enc_steps x beam)", (KM * d_.T + 8 * KM * 64) * sizeof(float)},
        {"attn_energy_kernel (decoder features, beam x 2 hidden_dim)", (KM * d_.A + 2 * (size_t)d_.A) * sizeof(float)},
        {"vocab_final_kernel (candidates + article words, max_enc_steps)", vocab_tc_ ? vtc_.final_smem(d_.T, vtc_.n_parts()) : 0},
        {"final_topk_kernel (fp32 vocabulary path: vocab_size bitmap + per-thread lists)",
         (size_t)((d_.V + 31) / 32) * 4 + (size_t)256 * K2M * 8 + 128}};
    for (auto& n : need)
      if (n.bytes > (size_t)optin) {
        *code = PGEN_ERR_UNSUPPORTED;
        return std::string(n.what) + " needs " + std::to_string(n.bytes) + " bytes of shared memory per block, this GPU allows " +
               std::to_string(optin) + ": lower max_enc_steps / beam_size";
      }
  }
  CK(cudaDeviceSynchronize());
  *code = PGEN_OK;
  return "";
}

// Kernel-only timing for the roofline: the attention pass and the tensor-core pass of the last decoded batch, each
// replayed `reps` times from a graph (both are idempotent on the buffers the batch left behind) between two CUDA events
// on this engine's stream -- no launch gaps.  Between two replays the graph overwrites a 256 MB scratch buffer, twice
// the L2: back to back the attention pass would find its 105 MB of encoder states / features, and the tensor-core
// pass its 51 MB of weights, in the 126 MB L2, which they never do in a decode step.  The same graph with the
// flushes alone is timed too and subtracted.  Results land in the profile fields of the stats (sums over `reps`
// launches, prof_steps = reps).  Re-initialises the beam state: call it after the batch's results have been taken.
std::string Engine::measure(int reps) {
  std::string err = sync();
  if (!err.empty()) return err;
  const int B = nB_, K = d_.K, R = B * K;
  if (B <= 0) return "";
  launch_init_beam(d_, B, lens_, bb_, st_);                    // every article active again
  const size_t flush_bytes = (size_t)256 << 20;
  if (!flush_) {
    CK(cudaMalloc((void**)&flush_, flush_bytes));
    allocs_.push_back(flush_);
  }
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  auto timed = [&](const std::function<void()>& body, double* total_ms) -> std::string {
    if (body) body();                                          // lazy function attributes outside the capture
    cudaGraph_t g = nullptr; cudaGraphExec_t ge = nullptr;
    if (cudaStreamBeginCapture(st_, cudaStreamCaptureModeRelaxed) != cudaSuccess) return "measure: capture failed";
    for (int i = 0; i < reps; ++i) {
      cudaMemsetAsync(flush_, i & 0xff, flush_bytes, st_);
      if (body) body();
    }
    if (cudaStreamEndCapture(st_, &g) != cudaSuccess || !g || cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) {
      if (g) cudaGraphDestroy(g);
      return "measure: graph instantiation failed";
    }
    cudaGraphDestroy(g);
    cudaGraphLaunch(ge, st_);                                  // warm-up replay
    cudaEventRecord(e0, st_);
    cudaGraphLaunch(ge, st_);
    cudaEventRecord(e1, st_);
    cudaStreamSynchronize(st_);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaGraphExecDestroy(ge);
    *total_ms = ms;
    return "";
  };
  double ms_flush = 0, ms_att = 0, ms_voc = 0;
  err = timed(nullptr, &ms_flush);
  if (err.empty()) err = timed([&] { attention(B, cov_[0], attn_[0], state_[0]); }, &ms_att);
  if (err.empty() && vocab_tc_)
    err = timed([&] { vtc_.launch(R, w_.proj_v, part_m_, part_s_, gmax_, gstride_, smax_, sstride_, nullptr, st_); }, &ms_voc);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (!err.empty()) return err;
  stats.ms_attention = std::max(0.0, ms_att - ms_flush); stats.ms_vocab = vocab_tc_ ? std::max(0.0, ms_voc - ms_flush) : 0.0;
  stats.ms_beam = 0; stats.prof_steps = (uint64_t)reps;
  return check("measure");
}

std::string Engine::set_option(const std::string& kv) {
  const size_t eq = kv.find('=');
  if (eq == std::string::npos) return "option must be key=value";
  const std::string k = kv.substr(0, eq), v = kv.substr(eq + 1);
  // captured step graphs bake the kernel variants in: drop them, they are re-captured on the next batch
  if (!graphs_.empty()) {
    CK(cudaSetDevice(device_));
    CK(cudaStreamSynchronize(st_));
    for (auto& g : graphs_) cudaGraphExecDestroy(g.second);
    graphs_.clear();
  }
  if (k == "graph") { use_graph_ = v != "0"; return ""; }
  if (k == "graph_steps") { graph_steps_ = std::max(0, atoi(v.c_str())); return ""; }
  if (k == "check_every") { check_every_ = std::max(0, atoi(v.c_str())); return ""; }
  if (k == "features") { feat_tc_ = (v != "simt") && g3_.ready; return ""; }
  if (k == "encoder") { enc_persistent_ = (v != "steps"); return ""; }
  if (k == "decoder") {
    if (v == "split") { dec_fused_ = false; return ""; }
    if (v == "fused") {
      if (!dcell_.ready) return "fused decoder cell needs hidden_dim 256 and emb_dim 128";
      dec_fused_ = true;
      return "";
    }
    return "decoder must be fused or split";
  }
  if (k == "attention") {
    if (v == "split") { attn_fused_ = false; return ""; }
    if (v == "fused") {
      if (!att_cnt_) return "one-launch attention needs hidden_dim 256 and beam_size 4 or 8";
      attn_fused_ = true;
      return "";
    }
    return "attention must be fused or split";
  }
  if (k == "vocab") {
    if (v == "simt") { vocab_tc_ = false; return ""; }
    if (v == "tc") {
      if (!vtc_.ready) return "tensor-core vocabulary path unavailable: " + vtc_why_;
      vocab_tc_ = true;
      return "";
    }
    return "vocab must be simt or tc";
  }
  if (k == "vocab_guard") {
    if (v == "auto") vocab_guard_ = 0; else if (v == "widen") vocab_guard_ = 1; else if (v == "dense") vocab_guard_ = 2;
    else if (v == "off") vocab_guard_ = 3; else return "vocab_guard must be auto, widen, dense or off";
    return "";
  }
  if (k == "measure") return ran_ ? measure(std::max(1, atoi(v.c_str()))) : std::string();
  if (k == "profile") {
    profile_ = atoi(v.c_str()) != 0;
    profile_first_ = std::max(0, atoi(v.c_str()));        // N > 1: the sums cover the first N steps only
    if (profile_ && pev_.empty()) {
      pev_.resize((size_t)5 * d_.S);
      for (auto& e : pev_) CK(cudaEventCreate(&e));
    }
    return "";
  }
  return "unknown option '" + k + "'";
}

// --------------------------------------------------------------------------------------
// Stage: H2D
// --------------------------------------------------------------------------------------
std::string Engine::upload(int n, const int32_t* enc_ids, const int32_t* ext_ids, int ld, const int32_t* lens,
                           const int32_t* n_oovs) {
  if (n < 0 || n > Bcap_) return "n_articles exceeds articles_per_batch";
  where_ = 1;
  CK(cudaSetDevice(device_));
  const int T = d_.T;
  nB_ = n;
  max_len_ = 0;
  std::vector<uint64_t> keys;                          // (extended id << 32) | position: sorting groups ids, positions ascending
  for (int b = 0; b < n; ++b) {
    const int len = lens[b];
    if (len < 0 || len > T || len > ld) return "enc_lens[" + std::to_string(b) + "] outside 0..max_enc_steps";
    const int no = n_oovs ? n_oovs[b] : 0;
    h_lens_[b] = len;
    h_noov_[b] = no;
    max_len_ = std::max(max_len_, len);
    int32_t* hi = h_ids_ + (size_t)b * T;
    int32_t* he = h_ext_ + (size_t)b * T;
    keys.resize((size_t)len);
    const int vext = d_.V + no;
    for (int t = 0; t < len; ++t) {
      const int id = enc_ids[(size_t)b * ld + t], ex = ext_ids[(size_t)b * ld + t];
      if (id < 0 || id >= d_.V) return "enc_ids out of range";
      if (ex < 0 || ex >= vext) return "enc_ext_ids outside the extended vocabulary";
      hi[t] = id; he[t] = ex;
      keys[(size_t)t] = ((uint64_t)(uint32_t)ex << 32) | (uint32_t)t;
    }
    for (int t = len; t < T; ++t) { hi[t] = 0; he[t] = 0; }
    // unique extended ids (ascending) with their positions in ascending t: the order in which
    // tf.scatter_nd accumulates duplicates (model.py:170-176)
    std::sort(keys.begin(), keys.end());
    int32_t* hu = h_uniq_ + (size_t)b * T;
    int32_t* hs = h_pstart_ + (size_t)b * (T + 1);
    int32_t* hp = h_plist_ + (size_t)b * T;
    int nu = 0;
    for (int q = 0; q < len; ++q) {
      const int ex = (int)(keys[(size_t)q] >> 32);
      if (q == 0 || ex != (int)(keys[(size_t)q - 1] >> 32)) { hu[nu] = ex; hs[nu] = q; ++nu; }
      hp[q] = (int)(uint32_t)keys[(size_t)q];
    }
    hs[nu] = len;
    h_ucnt_[b] = nu;
  }
  CK(cudaEventRecord(ev_[0], st_));
  const size_t nT = (size_t)n * T;
  if (n > 0) {
    CK(cudaMemcpyAsync(ids_, h_ids_, nT * 4, cudaMemcpyHostToDevice, st_));
    CK(cudaMemcpyAsync(ext_, h_ext_, nT * 4, cudaMemcpyHostToDevice, st_));
    CK(cudaMemcpyAsync(lens_, h_lens_, (size_t)n * 4, cudaMemcpyHostToDevice, st_));
    CK(cudaMemcpyAsync(noov_, h_noov_, (size_t)n * 4, cudaMemcpyHostToDevice, st_));
    CK(cudaMemcpyAsync(uniq_, h_uniq_, nT * 4, cudaMemcpyHostToDevice, st_));
    CK(cudaMemcpyAsync(ucnt_, h_ucnt_, (size_t)n * 4, cudaMemcpyHostToDevice, st_));
    CK(cudaMemcpyAsync(pstart_, h_pstart_, (size_t)n * (T + 1) * 4, cudaMemcpyHostToDevice, st_));
    CK(cudaMemcpyAsync(plist_, h_plist_, nT * 4, cudaMemcpyHostToDevice, st_));
  }
  CK(cudaEventRecord(ev_[1], st_));
  staged_ = true; ran_ = false; synced_ = false;
  return "";
}

// --------------------------------------------------------------------------------------
// Encoder: embedding gather fused into the input projection, BiLSTM, state reduce, W_h features
// --------------------------------------------------------------------------------------
void Engine::encoder() {
  const int B = nB_, T = d_.T, H = d_.H, A = d_.A, E = d_.E, K = d_.K;
  uint64_t& nl = stats.kernel_launches;
  cudaMemsetAsync(enc_states_, 0, (size_t)B * T * A * sizeof(float), st_);   // outputs past enc_len are zero
  cudaMemsetAsync(h_enc_[0], 0, (size_t)2 * B * H * sizeof(float), st_);
  cudaMemsetAsync(c_enc_, 0, (size_t)2 * B * H * sizeof(float), st_);
  const bool table = zx_tab_ && H == 256 && enc_persistent_;
  for (int dir = 0; dir < 2 && !table; ++dir) {        // zx = emb[ids] . Kx + bias   (model.py:213, LSTMCell)
    GemmArgs g;
    g.A0 = w_.emb; g.lda0 = E; g.k0 = E; g.gather0 = ids_;
    g.B = w_.enc_kxT[dir]; g.ldb = E; g.bias = w_.enc_b[dir];
    g.C = zx_ + (size_t)dir * B * T * 4 * H; g.ldc = 4 * H; g.M = B * T; g.N = 4 * H;
    nl += launch_gemm_nt(g, st_);
  }
  const float* hfin;
  cudaEventRecord(eev_[0], st_);
  if (H == 256 && enc_persistent_) {                   // model.py:90-93, one launch
    nl += launch_enc_lstm_persistent(d_, B, lens_, table ? zx_tab_ : zx_, table ? ids_ : nullptr, w_.enc_kh[0], h_enc_[1], c_enc_,
                                     enc_states_, st_);
    hfin = h_enc_[1];
  } else {
    for (int s = 0; s < max_len_; ++s)                 // model.py:90-93, one launch per time step
      nl += launch_enc_lstm_step(d_, B, s, lens_, zx_, w_.enc_kh[0], h_enc_[s & 1], h_enc_[(s + 1) & 1], c_enc_,
                                 enc_states_, st_);
    hfin = h_enc_[max_len_ & 1];
  }
  cudaEventRecord(eev_[1], st_);
  {                                                    // model.py:117-120 -> row b*K of the "old" state
    GemmArgs g;
    g.A0 = c_enc_; g.lda0 = H; g.k0 = H; g.A1 = c_enc_ + (size_t)B * H; g.lda1 = H; g.k1 = H;
    g.B = w_.red_wcT; g.ldb = A; g.bias = w_.red_bc; g.act = 1;
    g.C = state_[0] + A; g.ldc = K * d_.SW; g.M = B; g.N = H; g.skinny = 1;    // few rows, long K: split-K tile kernel
    nl += launch_gemm_nt(g, st_);
    g.A0 = hfin; g.A1 = hfin + (size_t)B * H; g.B = w_.red_whT; g.bias = w_.red_bh; g.C = state_[0] + A + H;
    nl += launch_gemm_nt(g, st_);
  }
  {                                                    // attention_decoder.py:67, once per article
    cudaEventRecord(eev_[2], st_);
    if (feat_tc_) {
      nl += launch_split_tf32(enc_states_, es_hi_, es_lo_, (size_t)B * T * A, st_);
      nl += g3_.launch(B * T, enc_feat_, A, st_);
    } else {
      GemmArgs g;
      g.A0 = enc_states_; g.lda0 = A; g.k0 = A; g.B = w_.WhT; g.ldb = A;
      g.C = enc_feat_; g.ldc = A; g.M = B * T; g.N = A;
      nl += launch_gemm_nt(g, st_);
    }
    cudaEventRecord(eev_[3], st_);
  }
}

// The attention pass of a decode step over all K rows of every live article (attention_decoder.py:156-158).
int Engine::attention(int B, const float* cov, float* attn, float* state) {
  if (attn_fused_) {
    int nl = dpack_fresh_ ? 0 : launch_attn_pack_dfeat(d_, B, dfeat_, dpack_, bb_.active, st_);
    dpack_fresh_ = false;
    return nl + launch_attn_fused(d_, B, lens_, bb_.active, enc_feat_, enc_states_, dpack_, cov, w_.v, w_.w_c, e_, att_pctx_,
                                  att_segstat_, att_cnt_, attn, state, d_.SW, n_sms_, st_);
  }
  int nl = launch_attn_energy(d_, B, d_.K, lens_, bb_.active, enc_feat_, dfeat_, cov, w_.v, w_.w_c, e_, st_);
  return nl + launch_attn_context(d_, B, d_.K, lens_, bb_.active, enc_states_, e_, attn, state, st_);
}

// One decode step (beam_search.py:109-156) for buffer parity o (old) -> n (new).  Nothing in it depends on the step
// number on the host side (beam_step reads it from steps_done), so a captured step can be replayed.
int Engine::step_launch(int B, int o, int ps) {
  const int T = d_.T, H = d_.H, A = d_.A, E = d_.E, K = d_.K, R = B * K, SW = d_.SW;
  const int n = o ^ 1;
  int nl = 0;
    StepBufs sb;
    sb.lens = lens_; sb.n_oov = noov_; sb.active = bb_.active; sb.sel_parent = bb_.sel_parent; sb.sel_token = bb_.sel_token;
    sb.emb = w_.emb; sb.state_old = state_[o]; sb.state_new = state_[n]; sb.attn_old = attn_[o]; sb.attn_new = attn_[n];
    sb.cov_old = cov_[o]; sb.cov_new = cov_[n]; sb.xin = xin_; sb.hprev = hprev_; sb.cprev = cprev_;
    GemmArgs g;
    if (dec_fused_) {
      // gather by parent, coverage update, x, LSTM cell and attention features in one launch (dec_cell.cu)
      nl += dcell_.launch_cell(d_, B, sb, w_.x_b, w_.dec_b, w_.att_b, dfeat_, attn_fused_ ? dpack_ : nullptr, st_);
      dpack_fresh_ = attn_fused_;
    } else {
    nl += launch_gather_step(d_, B, sb, st_);
    // x = [emb(y); ctx] . Linear/Matrix + Bias                                  (attention_decoder.py:150)
    g = GemmArgs(); g.A0 = xin_; g.lda0 = E + A; g.k0 = E + A; g.B = w_.x_MT; g.ldb = E + A; g.bias = w_.x_b;
    g.C = state_[n] + A + 2 * H; g.ldc = SW; g.M = R; g.N = E; g.skinny = 1;
    nl += launch_gemm_nt(g, st_);
    // LSTM cell                                                                   (:153)
    g = GemmArgs(); g.A0 = state_[n] + A + 2 * H; g.lda0 = SW; g.k0 = E; g.A1 = hprev_; g.lda1 = H; g.k1 = H;
    g.B = w_.dec_kT; g.ldb = E + H; g.bias = w_.dec_b; g.C = z_; g.ldc = 4 * H; g.M = R; g.N = 4 * H; g.skinny = 1;
    nl += launch_gemm_nt(g, st_);
    nl += launch_lstm_pointwise(d_, R, z_, cprev_, state_[n], bb_.active, st_);
    // attention on the new state, coverage not updated                            (:156-158)
    g = GemmArgs(); g.A0 = state_[n] + A; g.lda0 = SW; g.k0 = 2 * H; g.B = w_.att_MT; g.ldb = 2 * H; g.bias = w_.att_b;
    g.C = dfeat_; g.ldc = A; g.M = R; g.N = A; g.skinny = 1;
    if (attn_fused_ && (K & 1) == 0) { g.C_pack = dpack_; g.pack_K = K; dpack_fresh_ = true; }   // + the layout the attention kernel copies
    nl += launch_gemm_nt(g, st_);
    }
    if (profile_) cudaEventRecord(pev_[5 * ps + 0], st_);
    nl += attention(B, cov_[n], attn_[n], state_[n]);
    if (profile_) cudaEventRecord(pev_[5 * ps + 1], st_);
    // p_gen and the exact logits of the article's words are side branches (stream st2_, a parallel branch of the
    // captured graph): they run beside the out projection and the tensor-core pass and join before the final kernel
    const bool fork = d_.pointer_gen && vocab_tc_ && !profile_;
    if (dec_fused_) {
      // output = [h; ctx] . AttnOutputProjection + b (:173) and p_gen (:166-167) in one launch
      nl += dcell_.launch_out(d_, B, bb_.active, state_[n], w_.out_b, w_.pg_w, w_.pg_b, out_, vocab_tc_ ? out_tf32_ : nullptr,
                              pgen_, st_);
    } else {
    if (fork) { cudaEventRecord(fev_[0], st_); cudaStreamWaitEvent(st2_, fev_[0], 0); }
    if (d_.pointer_gen) nl += launch_pgen(d_, R, state_[n], w_.pg_w, w_.pg_b, pgen_, bb_.active, fork ? st2_ : st_);   // :166-167
    // output = [h; ctx] . AttnOutputProjection + b                                (:173)
    g = GemmArgs(); g.A0 = state_[n] + A + H; g.lda0 = SW; g.k0 = H; g.A1 = state_[n]; g.lda1 = SW; g.k1 = A;
    g.B = w_.out_MT; g.ldb = H + A; g.bias = w_.out_b; g.C = out_; g.ldc = H; g.M = R; g.N = H; g.skinny = 1;
    if (vocab_tc_) { g.C_tf32 = out_tf32_; g.ldc_tf32 = H; }   // the tensor-core pass multiplies the tf32-rounded copy
    nl += launch_gemm_nt(g, st_);
    }
    if (vocab_tc_) {
      // vocabulary scores on the tensor cores, candidates + softmax statistics in the epilogue   (model.py:236-238)
      int nmb, nr;
      vtc_.plan(R, &nmb, &nr);
      MergeArgs ma;
      ma.out = out_; ma.wT = w_.proj_wT; ma.bias = w_.proj_v; ma.part_m = part_m_; ma.part_s = part_s_;
      ma.gmax = gmax_; ma.gstride = gstride_; ma.smax = smax_; ma.sstride = sstride_;
      ma.artl = artl_; ma.pgen = pgen_; ma.attn = attn_[n]; ma.uniq_ids = uniq_;
      ma.uniq_cnt = ucnt_; ma.pos_start = pstart_; ma.pos_list = plist_; ma.active = bb_.active; ma.lens = lens_;
      ma.topk_ids = topk_ids_; ma.topk_logp = topk_logp_; ma.K = K; ma.K2 = d_.K2; ma.G = vtc_.G; ma.NC = vtc_.G * 8; ma.H = H;
      ma.V = d_.V; ma.T = T; ma.R = R; ma.n_parts = vtc_.n_parts(); ma.pointer_gen = d_.pointer_gen;
      ma.wnorm_max = w_.proj_wnorm_max; ma.guard = vocab_guard_; ma.sel_stats = sel_stats_;
      if (profile_) cudaEventRecord(pev_[5 * ps + 2], st_);
      if (fork) { cudaEventRecord(fev_[1], st_); cudaStreamWaitEvent(st2_, fev_[1], 0); }      // out_ is complete
      if (fork) { nl += vtc_.launch_rescore(ma, B, st2_); cudaEventRecord(fev_[2], st2_); }   // exact fp32 logits of the article's words
      nl += vtc_.launch(R, w_.proj_v, part_m_, part_s_, gmax_, gstride_, smax_, sstride_, nullptr, st_);
      if (profile_) cudaEventRecord(pev_[5 * ps + 3], st_);
      if (fork) cudaStreamWaitEvent(st_, fev_[2], 0);
      else nl += vtc_.launch_rescore(ma, B, st_);
      nl += vtc_.launch_final(ma, st_);                      // candidates re-scored in fp32, copy mass, top-2K   (model.py:146-183, 284-285)
      if (profile_) cudaEventRecord(pev_[5 * ps + 4], st_);
    } else {
    // vocabulary scores                                                           (model.py:236)
    g = GemmArgs(); g.A0 = out_; g.lda0 = H; g.k0 = H; g.B = w_.proj_wT; g.ldb = H; g.bias = w_.proj_v;
    g.C = logits_; g.ldc = d_.V; g.M = R; g.N = d_.V;
    if (profile_) cudaEventRecord(pev_[5 * ps + 2], st_);
    nl += launch_gemm_nt(g, st_);
    if (profile_) cudaEventRecord(pev_[5 * ps + 3], st_);
    nl += launch_softmax_stats(d_, R, logits_, rowmax_, rowsum_, bb_.active, st_);
    FinalArgs fa;
    fa.logits = logits_; fa.rowmax = rowmax_; fa.rowsum = rowsum_; fa.pgen = pgen_; fa.attn = attn_[n];
    fa.uniq_ids = uniq_; fa.uniq_cnt = ucnt_; fa.pos_start = pstart_; fa.pos_list = plist_; fa.n_oov = noov_;
    fa.active = bb_.active; fa.topk_ids = topk_ids_; fa.topk_logp = topk_logp_;
    nl += launch_final_topk(d_, B, fa, st_);
    if (profile_) cudaEventRecord(pev_[5 * ps + 4], st_);
    }
    nl += launch_beam_step(d_, B, topk_ids_, topk_logp_, bb_, st_);
  return nl;
}

// `chunk` consecutive steps (even count) captured once per batch size into an executable graph: a kernel node costs
// ~0.6 us of launch overhead instead of the ~4 us (quantised) of a stream launch on this platform.
cudaGraphExec_t Engine::step_graph(int B, int chunk, bool capture) {
  if (graph_failed_) return nullptr;
  if (graph_chunk_ != chunk) {
    for (auto& kv : graphs_) cudaGraphExecDestroy(kv.second);
    graphs_.clear();
    graph_chunk_ = chunk;
  }
  auto it = graphs_.find(B);
  if (it != graphs_.end()) return it->second;
  if (!capture) return nullptr;
  cudaGraph_t g = nullptr;
  cudaGraphExec_t ge = nullptr;
  int nodes = 0;
  if (cudaStreamBeginCapture(st_, cudaStreamCaptureModeRelaxed) != cudaSuccess) { graph_failed_ = true; cudaGetLastError(); return nullptr; }
  for (int i = 0; i < chunk; ++i) nodes += step_launch(B, i & 1, 0);
  if (cudaStreamEndCapture(st_, &g) != cudaSuccess || !g || cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) {
    if (g) cudaGraphDestroy(g);
    graph_failed_ = true; cudaGetLastError();
    return nullptr;
  }
  cudaGraphDestroy(g);
  graph_nodes_ = nodes;
  graphs_[B] = ge;
  return ge;
}

// --------------------------------------------------------------------------------------
// Decoder: initial attention pass + step loop
// --------------------------------------------------------------------------------------
void Engine::decoder() {
  const int B = nB_, T = d_.T, H = d_.H, A = d_.A, K = d_.K, R = B * K, SW = d_.SW;
  uint64_t& nl = stats.kernel_launches;
  nl += launch_init_beam(d_, B, lens_, bb_, st_);
  if (att_cnt_) cudaMemsetAsync(att_cnt_, 0, (size_t)(B + 1) * sizeof(int), st_);   // segment counters of the one-launch attention
  if (d_.use_cov) cudaMemsetAsync(cov_[0], 0, (size_t)R * T * sizeof(float), st_);   // beam_search.py:105
  {  // attention_decoder.py:138-140: context of the initial state (row b*K only; all K hyps are identical)
    GemmArgs g;
    g.A0 = state_[0] + A; g.lda0 = K * SW; g.k0 = 2 * H; g.B = w_.att_MT; g.ldb = 2 * H; g.bias = w_.att_b;
    g.C = dfeat_; g.ldc = K * A; g.M = B; g.N = A; g.skinny = 1;
    nl += launch_gemm_nt(g, st_);
    nl += launch_attn_energy(d_, B, 1, lens_, bb_.active, enc_feat_, dfeat_, cov_[0], w_.v, w_.w_c, e_, st_);
    nl += launch_attn_context(d_, B, 1, lens_, bb_.active, enc_states_, e_, attn_[0], state_[0], st_);
  }
  int steps = 0;
  const bool graphs = use_graph_ && !profile_ && graph_steps_ >= 2;
  const int chunk = graphs ? (graph_steps_ & ~1) : 0;          // an even number of steps per graph: buffer parity restarts
  cudaGraphExec_t ge = graphs ? step_graph(B, chunk, false) : nullptr;
  if (graphs && !ge && chunk <= d_.S) {
    // first batch of this size: one chunk with plain launches (which also sets the kernels' lazy function attributes
    // outside any capture), then capture; a failed capture leaves plain launches for good
    where_ = 12;
    for (int i = 0; i < chunk; ++i) nl += step_launch(B, i & 1, i);
    steps = chunk;
    where_ = 13;
    ge = step_graph(B, chunk, true);
  }
  where_ = 15;
  while (ge && steps + chunk <= d_.S) {
    cudaGraphLaunch(ge, st_);
    nl += (uint64_t)graph_nodes_;
    steps += chunk;
    if (check_every_ > 0 && steps < d_.S && steps > d_.min_dec_steps) {
      cudaMemcpyAsync(h_nactive_, bb_.n_active, sizeof(int), cudaMemcpyDeviceToHost, st_);
      cudaStreamSynchronize(st_);
      if (*h_nactive_ == 0) { stats.decode_steps += steps; pev_used_ = 0; goto done; }
    }
  }
  for (int s = steps; s < d_.S; ++s) {
    nl += step_launch(B, s & 1, s);
    ++steps;
    if (check_every_ > 0 && s + 1 < d_.S && s + 1 > d_.min_dec_steps && (s + 1) % check_every_ == 0) {
      cudaMemcpyAsync(h_nactive_, bb_.n_active, sizeof(int), cudaMemcpyDeviceToHost, st_);
      cudaStreamSynchronize(st_);
      if (*h_nactive_ == 0) break;
    }
  }
  stats.decode_steps += steps;
  pev_used_ = profile_ ? steps : 0;
done:
  nl += launch_finalize(d_, B, lens_, bb_, out_tok_, out_len_, out_lp_, out_avg_, st_);
}

std::string Engine::run() {
  if (!staged_) return "pgen_stage_run before pgen_stage_upload";
  where_ = 10;
  CK(cudaSetDevice(device_));
  CK(cudaEventRecord(ev_[2], st_));
  if (nB_ > 0) encoder();
  where_ = 11;
  CK(cudaEventRecord(ev_[3], st_));
  if (nB_ > 0) decoder();
  where_ = 17;
  CK(cudaEventRecord(ev_[4], st_));
  // results to pinned host memory on the same stream: the batch is complete when ev_[6] has fired
  const size_t n = nB_, S = d_.S;
  CK(cudaEventRecord(ev_[5], st_));
  if (n > 0) {
    CK(cudaMemcpyAsync(h_out_tok_, out_tok_, n * S * 4, cudaMemcpyDeviceToHost, st_));
    CK(cudaMemcpyAsync(h_out_len_, out_len_, n * 4, cudaMemcpyDeviceToHost, st_));
    CK(cudaMemcpyAsync(h_out_lp_, out_lp_, n * S * 4, cudaMemcpyDeviceToHost, st_));
    CK(cudaMemcpyAsync(h_out_avg_, out_avg_, n * 8, cudaMemcpyDeviceToHost, st_));
    CK(cudaMemcpyAsync(h_steps_, bb_.steps_done, n * 4, cudaMemcpyDeviceToHost, st_));
    CK(cudaMemcpyAsync(h_rescnt_, bb_.res_count, n * 4, cudaMemcpyDeviceToHost, st_));
    CK(cudaMemcpyAsync(h_selstats_, sel_stats_, 2 * 4, cudaMemcpyDeviceToHost, st_));
  }
  CK(cudaEventRecord(ev_[6], st_));
  std::string err = check("decode launch");
  if (!err.empty()) return err;
  ran_ = true; synced_ = false; counted_ = false;
  ++stats.batches;
  where_ = 19;
  return "";
}

#ifdef PGEN_LIVE_VT
void live_read_vt(unsigned int* out16, cudaStream_t st);
#endif
#ifdef PGEN_LIVE_AF
void live_read_af(unsigned int* out16, cudaStream_t st);
#endif
#ifdef PGEN_LIVE_PE
void live_read_pe(unsigned int* out16, cudaStream_t st);
#endif
// entry / exit counters of the kernels that wait on mbarriers (zeros unless built with -DPGEN_LIVE): 3 x 16 values
void Engine::debug_live(long long* out48) {
  for (int i = 0; i < 48; ++i) out48[i] = 0;
#if defined(PGEN_LIVE_VT) || defined(PGEN_LIVE_AF) || defined(PGEN_LIVE_PE)
  if (cudaSetDevice(device_) != cudaSuccess) return;
  cudaStream_t s = nullptr;
  if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) return;
  unsigned int v[48] = {};
#ifdef PGEN_LIVE_VT
  live_read_vt(v, s);
#endif
#ifdef PGEN_LIVE_AF
  live_read_af(v + 16, s);
#endif
#ifdef PGEN_LIVE_PE
  live_read_pe(v + 32, s);
#endif
  cudaStreamSynchronize(s);
  cudaStreamDestroy(s);
  for (int i = 0; i < 48; ++i) out48[i] = v[i];
#endif
}

void Engine::debug_progress(long long* out3) {
  out3[0] = out3[1] = out3[2] = -1;
  if (cudaSetDevice(device_) != cudaSuccess) return;
  cudaStream_t s = nullptr;
  if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) return;
  int v[3] = {-1, -1, -1};
  cudaMemcpyAsync(&v[0], bb_.n_active, 4, cudaMemcpyDeviceToHost, s);
  cudaMemcpyAsync(&v[1], bb_.steps_done, 4, cudaMemcpyDeviceToHost, s);
  if (att_cnt_) cudaMemcpyAsync(&v[2], att_cnt_, 4, cudaMemcpyDeviceToHost, s);
  cudaStreamSynchronize(s);
  cudaStreamDestroy(s);
  for (int i = 0; i < 3; ++i) out3[i] = v[i];
}

bool Engine::done() {
  if (!ran_) return true;
  cudaSetDevice(device_);
  return cudaEventQuery(ev_[6]) == cudaSuccess;
}

std::string Engine::sync() {
  where_ = 20;
  CK(cudaSetDevice(device_));
  CK(cudaStreamSynchronize(st_));
  where_ = 21;
  if (ran_ && !synced_) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, ev_[0], ev_[1]) == cudaSuccess) stats.ms_h2d = ms;
    if (cudaEventElapsedTime(&ms, ev_[2], ev_[3]) == cudaSuccess) stats.ms_encoder = ms;
    if (cudaEventElapsedTime(&ms, ev_[3], ev_[4]) == cudaSuccess) stats.ms_decoder = ms;
    if (cudaEventElapsedTime(&ms, ev_[5], ev_[6]) == cudaSuccess) stats.ms_d2h = ms;
    if (nB_ > 0 && cudaEventElapsedTime(&ms, eev_[0], eev_[1]) == cudaSuccess) stats.ms_enc_recurrence = ms;
    if (nB_ > 0 && cudaEventElapsedTime(&ms, eev_[2], eev_[3]) == cudaSuccess) stats.ms_enc_features = ms;
    if (pev_used_ > 0) {
      stats.ms_attention = stats.ms_vocab = stats.ms_beam = 0;
      const int upto = profile_first_ > 1 ? std::min(profile_first_, pev_used_) : pev_used_;
      for (int s = 0; s < upto; ++s) {
        if (cudaEventElapsedTime(&ms, pev_[5 * s], pev_[5 * s + 1]) == cudaSuccess) stats.ms_attention += ms;
        if (cudaEventElapsedTime(&ms, pev_[5 * s + 2], pev_[5 * s + 3]) == cudaSuccess) stats.ms_vocab += ms;
        if (cudaEventElapsedTime(&ms, pev_[5 * s + 3], pev_[5 * s + 4]) == cudaSuccess) stats.ms_beam += ms;
      }
      stats.prof_steps = upto;
    }
    synced_ = true;
  }
  return check("decode");
}

std::string Engine::download(int expect_rows, int32_t* out_tokens, int32_t* out_lens, float* out_logp, double* out_avg) {
  if (!ran_) return "pgen_stage_download before pgen_stage_run";
  if (expect_rows >= 0 && expect_rows != nB_)
    return "the slot holds a batch of " + std::to_string(nB_) + " articles, the caller expects " + std::to_string(expect_rows);
  std::string err = sync();
  if (!err.empty()) return err;
  const size_t n = nB_, S = d_.S;
  if (!counted_) {
    for (size_t b = 0; b < n; ++b) {
      stats.article_steps += h_steps_[b];
      if (h_rescnt_[b] >= d_.K) ++stats.early_stops;
    }
    stats.selection_widened = (uint64_t)h_selstats_[0];
    stats.selection_fallbacks = (uint64_t)h_selstats_[1];
    counted_ = true;
  }
  if (out_tokens) memcpy(out_tokens, h_out_tok_, n * S * 4);
  if (out_lens) memcpy(out_lens, h_out_len_, n * 4);
  if (out_logp) memcpy(out_logp, h_out_lp_, n * S * 4);
  if (out_avg) memcpy(out_avg, h_out_avg_, n * 8);
  return "";
}

std::string Engine::vocab_logits_op(int rows, const float* out_dev, float* logits_dev, int use_tc) {
  if (rows <= 0 || rows > R_) return "rows must be in 1..articles_per_batch*beam_size";
  CK(cudaSetDevice(device_));
  const int H = d_.H, V = d_.V;
  if (use_tc) {
    if (!vtc_.ready) return "tensor-core vocabulary path unavailable: " + vtc_why_;
    launch_round_tf32(out_dev, out_tf32_, (size_t)rows * H, st_);
    vtc_.launch(rows, w_.proj_v, part_m_, part_s_, gmax_, gstride_, smax_, sstride_, logits_dev, st_);
  } else {
    GemmArgs g;
    g.A0 = out_dev; g.lda0 = H; g.k0 = H; g.B = w_.proj_wT; g.ldb = H; g.bias = w_.proj_v;
    g.C = logits_dev; g.ldc = V; g.M = rows; g.N = V;
    launch_gemm_nt(g, st_);
  }
  CK(cudaStreamSynchronize(st_));
  return check("pgen_debug_vocab_logits");
}

std::string Engine::attention_op(int n, const float* enc_feat, const float* enc_states, const int32_t* lens,
                                 const float* dfeat, const float* cov, float* attn, float* ctx) {
  if (n <= 0) return "";
  CK(cudaSetDevice(device_));
  const size_t rows = (size_t)n * d_.K;
  const bool fused = attn_fused_ && att_cnt_ && n <= attn_fused_max_articles();
  // scratch of the call: energies, "article is live" flags, and for the one-launch kernel its per-segment partials
  const size_t n_pctx = fused ? attn_fused_pctx_floats(d_, n) : 0, n_seg = fused ? attn_fused_segstat_floats(d_, n) : 0;
  const size_t n_pack = fused ? rows * d_.A : 0;
  float* e = nullptr;
  int* act = nullptr;
  CK(cudaMalloc((void**)&e, (rows * d_.T + n_pctx + n_seg + n_pack) * sizeof(float)));
  if (cudaMalloc((void**)&act, ((size_t)2 * n + 1) * sizeof(int)) != cudaSuccess) { cudaFree(e); return "cudaMalloc failed"; }
  std::vector<int> ones((size_t)2 * n + 1, 1);
  std::fill(ones.begin() + n, ones.end(), 0);                     // [n] live flags, [n] segment counters, ticket counter
  cudaMemcpyAsync(act, ones.data(), ((size_t)2 * n + 1) * sizeof(int), cudaMemcpyHostToDevice, st_);
  Dims d = d_;
  d.SW = d.A;                      // ctx rows are dense [rows, A] here
  if (fused) {
    float* pack = e + rows * d_.T + n_pctx + n_seg;
    stats.kernel_launches += launch_attn_pack_dfeat(d, n, dfeat, pack, nullptr, st_);
    stats.kernel_launches += launch_attn_fused(d, n, lens, act, enc_feat, enc_states, pack, cov, w_.v, w_.w_c, e,
                                               e + rows * d_.T, e + rows * d_.T + n_pctx, act + n, attn, ctx, d.A, n_sms_, st_);
  } else {
    stats.kernel_launches += launch_attn_energy(d, n, d.K, lens, act, enc_feat, dfeat, cov, w_.v, w_.w_c, e, st_);
    stats.kernel_launches += launch_attn_context(d, n, d.K, lens, act, enc_states, e, attn, ctx, st_);
  }
  cudaError_t err = cudaStreamSynchronize(st_);
  cudaFree(e);
  cudaFree(act);
  if (err != cudaSuccess) return std::string("pgen_attention: ") + cudaGetErrorString(err);
  return check("pgen_attention");
}

std::string Engine::fetch(const std::string& name, void* out, size_t cap_bytes, size_t* n) {
  if (!ran_) return "pgen_debug_fetch before pgen_stage_run";
  CK(cudaSetDevice(device_));
  CK(cudaStreamSynchronize(st_));
  const size_t B = nB_, T = d_.T, A = d_.A, H = d_.H, S = d_.S, K = d_.K;
  const void* src = nullptr;
  size_t cnt = 0, esz = 4;
  if (name == "enc_states") { src = enc_states_; cnt = B * T * A; }
  else if (name == "enc_feat") { src = enc_feat_; cnt = B * T * A; }
  else if (name == "trace_tokens") { src = bb_.hist_tok; cnt = B * S * K; }
  else if (name == "trace_parents") { src = bb_.hist_par; cnt = B * S * K; }
  else if (name == "trace_logp") { src = bb_.hist_logp; cnt = B * S * K; }
  else if (name == "steps_done") { src = bb_.steps_done; cnt = B; }
  else if (name == "dec_c0" || name == "dec_h0") {
    // the initial state is overwritten by step 1 (double buffering): only valid with max_dec_steps == 1
    cnt = B * H;
    if (cap_bytes < cnt * 4) return "buffer too small";
    const float* base = state_[0] + A + (name == "dec_h0" ? H : 0);
    CK(cudaMemcpy2D(out, H * 4, base, (size_t)K * d_.SW * 4, H * 4, B, cudaMemcpyDeviceToHost));
    *n = cnt;
    return "";
  } else return "unknown debug buffer '" + name + "'";
  if (cap_bytes < cnt * esz) return "buffer too small";
  if (cnt) CK(cudaMemcpy(out, src, cnt * esz, cudaMemcpyDeviceToHost));
  *n = cnt;
  return "";
}

}  // namespace pgn
