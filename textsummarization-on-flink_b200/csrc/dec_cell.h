// Fused decoder cell of a beam-search step and the output projection + p_gen that follows the attention pass
// (hidden_dim 256, emb_dim 128 only).  See dec_cell.cu.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <string>

#include "kernels.h"

namespace pgn {

class DecCell {
 public:
  // The weight pointers are the engine's [N_out, K_in] device copies; init() makes the permuted copies whose slices
  // each CTA of a cluster streams contiguously, and the TMA descriptors.  `alloc(n_floats)` hands out device memory
  // that lives as long as the weights.  Returns "" or why the kernels cannot be used.
  template <class Alloc>
  std::string init(const Dims& d, const float* x_MT, const float* dec_kT, const float* att_MT, const float* out_MT, Alloc alloc,
                   cudaStream_t st) {
    if (d.H != 256 || d.E != 128) return "fused decoder cell needs hidden_dim 256 and emb_dim 128";
    float* k2 = alloc((size_t)4 * d.H * (d.E + d.H));
    float* k3 = alloc((size_t)d.A * d.A);
    if (!k2 || !k3) return "cudaMalloc failed for the permuted decoder weights";
    return setup(x_MT, dec_kT, att_MT, out_MT, k2, k3, st);
  }
  // the other engines of a handle read the same weights: same descriptors
  void share(const DecCell& o) { *this = o; }
  int launch_cell(const Dims& d, int B, const StepBufs& b, const float* x_b, const float* dec_b, const float* att_b,
                  float* dfeat, float* dpack, cudaStream_t st) const;
  int launch_out(const Dims& d, int B, const int* active, const float* state, const float* out_b, const float* pg_w,
                 const float* pg_b, float* out, float* out_tf32, float* pgen, cudaStream_t st) const;
  bool ready = false;

 private:
  std::string setup(const float* x_MT, const float* dec_kT, const float* att_MT, const float* out_MT, float* k2, float* k3,
                    cudaStream_t st);
  CUtensorMap tmW1, tmW2, tmW3, tmWo;
};

}  // namespace pgn
