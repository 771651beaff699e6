// Micro-benchmark of the per-step decoder GEMMs (not part of the product): nvcc -O3 -gencode arch=compute_100a,code=sm_100a
// NOTE: back-to-back stream launches are quantised to ~2 us ticks on this platform (tick_bench.cu): compare variants
// inside a CUDA graph or with ncu, not with this loop alone.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../textsummarization-on-flink_b200/csrc/kernels.cu"
using namespace pgn;
int main() {
  const int R = 256, H = 256, E = 128, A = 512;
  struct Case { const char* name; int N, k0, k1; } cases[] = {{"x", E, E + A, 0}, {"lstm", 4 * H, E, H}, {"dfeat", A, 2 * H, 0}, {"out", H, H, A}};
  float *a0, *a1, *b, *c, *bias;
  cudaMalloc(&a0, R * 1024 * 4); cudaMalloc(&a1, R * 1024 * 4); cudaMalloc(&b, 1024 * 1024 * 4); cudaMalloc(&c, R * 1024 * 4); cudaMalloc(&bias, 1024 * 4);
  cudaMemset(a0, 0, R * 1024 * 4); cudaMemset(a1, 0, R * 1024 * 4); cudaMemset(b, 0, 1024 * 1024 * 4); cudaMemset(bias, 0, 4096);
  cudaStream_t st; cudaStreamCreate(&st);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int mode = 0; mode < 1; ++mode) {
    for (auto& cs : cases) {
      GemmArgs g; g.A0 = a0; g.lda0 = 1024; g.k0 = cs.k0; g.A1 = a1; g.lda1 = 1024; g.k1 = cs.k1; g.B = b; g.ldb = cs.k0 + cs.k1;
      g.bias = bias; g.C = c; g.ldc = 1024; g.M = R; g.N = cs.N; g.skinny = 1;
      for (int i = 0; i < 20; ++i) launch_gemm_nt(g, st);
      cudaEventRecord(e0, st);
      for (int i = 0; i < 200; ++i) launch_gemm_nt(g, st);
      cudaEventRecord(e1, st); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      printf("%s: %.2f us per launch (%s)\n", cs.name, ms * 1000 / 200, cudaGetErrorString(cudaGetLastError()));
    }
  }
  for (int M : {32, 128, 256}) for (int N : {128, 512, 1024}) for (int K : {256, 512, 1024}) {
    GemmArgs g; g.A0 = a0; g.lda0 = 1024; g.k0 = K; g.B = b; g.ldb = K; g.bias = bias; g.C = c; g.ldc = 1024; g.M = M; g.N = N; g.skinny = 1;
    for (int i = 0; i < 20; ++i) launch_gemm_nt(g, st);
    cudaEventRecord(e0, st);
    for (int i = 0; i < 200; ++i) launch_gemm_nt(g, st);
    cudaEventRecord(e1, st); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("M %d N %d K %d: %.2f us\n", M, N, K, ms * 1000 / 200);
  }
  return 0;
}
