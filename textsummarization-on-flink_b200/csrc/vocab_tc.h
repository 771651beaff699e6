// Tensor-core (tcgen05, tf32) vocabulary projection fused with softmax statistics and candidate-group maxima, the
// per-row group selection + exact re-scoring, and the merge kernel.  See vocab_tc.cu.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <string>

namespace pgn {

struct MergeArgs {
  const float* out;        // [R, H] unrounded decoder outputs
  const float* wT;         // [V, H] unrounded output_projection/w (transposed)
  const float* bias;       // [V]
  const float* part_m; const float* part_s;      // [n_parts, R]   log2 domain (max, sum exp2) of the tf32 logits, per 128 words
  const float4* gmax; int gstride;               // [R, gstride] per row: maxima of the 4 groups of 8 words of each 32-word chunk
  const float* smax; int sstride;                // [R, sstride] per row: maxima of the 128-word supergroups
  float* artl;                                   // [R, T] exact logits of the article's distinct words
  const float* pgen; const float* attn;          // [R], [R, T]
  const int* uniq_ids; const int* uniq_cnt; const int* pos_start; const int* pos_list;
  const int* active; const int* lens;
  int* topk_ids; float* topk_logp;               // [R, 2K]
  int K, K2, G, NC, H, V, T, R, n_parts, pointer_gen;
  float wnorm_max = 0.f;                         // max over the vocabulary of |w_w|_2 (error bound of the tf32 logits)
  int guard = 0;                                 // 0 auto, 1 always widen, 2 always dense, 3 off (tests / A-B runs)
  int* sel_stats = nullptr;                      // [2] rows whose candidate set was widened / scored densely
};

class VocabTC {
 public:
  // wT_tf32 [V, H] and out_tf32 [Rcap (multiple of 128), H] are tf32-rounded device copies whose
  // addresses stay fixed (they are baked into the TMA descriptors).
  std::string init(const float* wT_tf32, int V, int H, const float* out_tf32, int Rcap, int K2, int n_sms);
  void plan(int R, int* n_mpairs, int* n_ranges) const;
  int n_parts() const;            // (max, sum) statistics per row: one per half tile of 128 words, whatever the grid
  int launch(int R, const float* bias, float* part_m, float* part_s, float4* gmax, int gstride, float* smax, int sstride, float* logits_dbg,
             cudaStream_t st) const;
  int launch_rescore(const MergeArgs& a, int B, cudaStream_t st) const;     // article words
  int launch_final(const MergeArgs& a, cudaStream_t st) const;              // candidates, final distribution, top-2K
  size_t smem_bytes() const;
  size_t final_smem(int T, int n_parts) const;
  bool ready = false;
  int G = 16;                       // candidate groups (of 8 words) re-scored per row = 2K + 8

 private:
  CUtensorMap tmA, tmB;
  int V = 0, H = 0, K2 = 0, KB = 0, n_tiles = 0, sms = 148;
};

// C[rows, N] = A[rows, K] . B[N, K]^T on the tensor cores with split operands (3xTF32, ~1e-5 absolute on O(1) sums);
// the operands are given as their
// tf32 hi / lo parts (launch_split_tf32), at fixed device addresses baked into the TMA descriptors.
class GemmTC3 {
 public:
  std::string init(const float* a_hi, const float* a_lo, int M, const float* b_hi, const float* b_lo, int N, int K, int n_sms);
  int launch(int rows, float* C, int ldc, cudaStream_t st) const;
  bool ready = false;

 private:
  CUtensorMap tmAh, tmAl, tmBh, tmBl;
  int M = 0, N = 0, K = 0, sms = 148;
};

int launch_round_tf32(const float* in, float* out, size_t n, cudaStream_t st);
int launch_split_tf32(const float* in, float* hi, float* lo, size_t n, cudaStream_t st);   // n % 4 == 0

}  // namespace pgn
