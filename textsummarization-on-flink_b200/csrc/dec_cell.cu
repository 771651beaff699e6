// The decoder cell of one beam-search step in ONE launch (attention_decoder.py:150-158, beam_search.py:109-124):
//
//   gather by parent      row r continues hypothesis sel_parent[r] with token sel_token[r] (OOV -> [UNK])
//   cov += attn           coverage of the parent + its attention (:114), for the row's own positions
//   x  = [emb(y); ctx] . Linear/Matrix + Bias                                   640 -> 128     (:150)
//   z  = [x; h] . lstm_cell/kernel + bias,  (c, h) = LSTM gates(z, c)           384 -> 4 x 256 (:153)
//   d  = [c; h] . Attention/Linear/Matrix + Bias                                 512 -> 512     (:93, :156-158)
//
// replacing gather_step + three split-K GEMM launches + lstm_pointwise (each 10-13 us of latency for a few MFLOP).
// A cluster of 8 CTAs owns a block of 16 decoder rows; the three products depend on each other, but only through the
// 16 rows of the block, so nothing is exchanged outside the cluster.  CTA c computes columns [16c, +16) of x, the
// four gates of hidden units [32c, +32) and columns [64c, +64) of d for the 16 rows.  Its slices of the three weight
// matrices (360 KB) stream through a ring of four 16 KB shared-memory stages filled by a producer warp with TMA tensor
// copies (128-byte swizzle, so that "lane = output column" reads are conflict-free) completing on mbarriers -- the
// stream runs ahead across the stages of the computation, the weights do not depend on the activations.  The
// activations stay in shared memory as [row][k]; x and (c, h) are exchanged between the 8 CTAs with ONE
// shared::cta -> shared::cluster bulk copy per destination, completing on the receiver's mbarrier.  All arithmetic is
// fp32 FMA (these values feed the recurrent state for 100 steps): packed FFMA2 whose halves are the even and the odd
// k of one output, 8 rows x up to 4 columns of outputs per thread, k split over the warps and combined through shared
// memory in fixed order.
//
// dec_out_kernel is the same machinery for output = [h; ctx] . AttnOutputProjection + b (:173) fused with
// p_gen = sigmoid([ctx; c; h; x] . w + b) (:166-167), after the attention pass.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

#include "dec_cell.h"

namespace pgn {

namespace {

constexpr int DC_H = 256, DC_E = 128, DC_A = 512, DC_SW = DC_A + 2 * DC_H + DC_E;
constexpr int DC_RB = 16;                            // decoder rows per cluster
constexpr int DC_CL = 8;                             // CTAs per cluster
constexpr int DC_NC = 512, DC_NT = DC_NC + 64;       // compute threads (16 warps: four per scheduler, so that one warp's
                                                     // dependent FFMA2 pairs and shared-memory latencies hide behind the others), + one
                                                     // producer warp, + one warp for the coverage update
constexpr int DC_K1 = DC_E + DC_A, DC_K2 = DC_E + DC_H, DC_K3 = 2 * DC_H;      // 640, 384, 512
constexpr int DC_N1 = DC_E / DC_CL, DC_N2 = 4 * DC_H / DC_CL, DC_N3 = DC_A / DC_CL;   // 16, 128, 64 columns per CTA
constexpr int DC_STAGE = 16384, DC_STAGES = 4;
constexpr int DC_CH1 = DC_K1 / 128, DC_CH2 = DC_K2 / 32, DC_CH3 = DC_K3 / 64;  // 5, 12, 8 chunks
constexpr float kS = 2.885390081777927f;             // 2 log2(e), see attn_fused.cu

// shared memory (bytes)
constexpr int OFF_RING = 0;
constexpr int OFF_A1 = OFF_RING + DC_STAGES * DC_STAGE;          // [16][640]  emb(y) | ctx of the parent
constexpr int OFF_A2X = OFF_A1 + DC_RB * DC_K1 * 4;              // [8 src][16][16]  x
constexpr int OFF_A2H = OFF_A2X + DC_RB * DC_E * 4;              // [16][256]  h of the parent
constexpr int OFF_A3 = OFF_A2H + DC_RB * DC_H * 4;               // [8 src][16][64]  new (c | h), 32 + 32 units per source
constexpr int OFF_RED = OFF_A3 + DC_RB * DC_K3 * 4;              // [kg][16][cols] partial sums, <= 32 KB
constexpr int OFF_CPREV = OFF_RED + 32768;                       // [16][32]  c of the parent, this CTA's units
constexpr int OFF_XOUT = OFF_CPREV + DC_RB * 32 * 4;             // [16][16]  this CTA's x, source of the exchange
constexpr int OFF_CHOUT = OFF_XOUT + DC_RB * DC_N1 * 4;          // [16][64]  this CTA's (c | h)
constexpr int OFF_BAR = OFF_CHOUT + DC_RB * 64 * 4;
constexpr int OFF_ROWOK = OFF_BAR + 128;                         // [16] row exists and its article is still being decoded
constexpr int DC_SMEM = OFF_ROWOK + 64;

__device__ __forceinline__ uint32_t dc_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void dc_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = dc_u32(bar);
  uint32_t ok = 0, spins = 0;
  while (!ok) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    if (!ok && ++spins > (1u << 26)) __trap();       // a protocol bug becomes an error, not a hang
  }
}
__device__ __forceinline__ void dc_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dc_u32(bar)), "r"(count));
}
__device__ __forceinline__ void dc_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(dc_u32(bar)) : "memory");
}
__device__ __forceinline__ void dc_expect(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(dc_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void dc_bulk(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dc_u32(dst)), "l"(src), "r"(bytes), "r"(dc_u32(bar)) : "memory");
}
// one box of a weight matrix: 32 consecutive k (128 B, swizzled) of `rows` consecutive output columns
__device__ __forceinline__ void dc_tma(void* dst, const CUtensorMap* map, int k0, int row0, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dc_u32(dst)), "l"(map), "r"(dc_u32(bar)), "r"(k0), "r"(row0) : "memory");
}
__device__ __forceinline__ uint32_t dc_mapa(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void dc_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void dc_sync_compute() { asm volatile("bar.sync 1, %0;" ::"n"(DC_NC) : "memory"); }
__device__ __forceinline__ unsigned long long dc_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ float dc_sum2(unsigned long long v) {
  float lo, hi;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
  return lo + hi;
}
__device__ __forceinline__ float dc_sigmoid(float x) { return 1.f / (1.f + expf(-x)); }

// Four consecutive k of NCOL output columns x 8 rows: acc[c][r] (+)= A[r][k..k+3] . W[col_c][k..k+3], the two halves of
// every packed FMA being the even and the odd k.  `arow` points at A[first row][k] (rows `apitch` floats apart, every
// lane reads the same addresses: broadcast); `wtile` is the swizzled weight box, jj the 16-byte chunk of the step.
template <int NCOL>
__device__ __forceinline__ void dc_step(unsigned long long (&acc)[NCOL][8], const float* arow, int apitch,
                                        const unsigned char* wtile, const int (&col)[NCOL], int jj) {
  ulonglong2 av[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) av[r] = *reinterpret_cast<const ulonglong2*>(arow + r * apitch);
#pragma unroll
  for (int c = 0; c < NCOL; ++c) {
    const ulonglong2 wv = *reinterpret_cast<const ulonglong2*>(wtile + col[c] * 128 + ((jj ^ (col[c] & 7)) << 4));
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      acc[c][r] = dc_fma2(av[r].x, wv.x, acc[c][r]);
      acc[c][r] = dc_fma2(av[r].y, wv.y, acc[c][r]);
    }
  }
}

}  // namespace

struct DcArgs {
  const int* lens; const int* active; const int* sel_parent; const int* sel_token;
  const float* emb;
  const float* state_old; float* state_new;          // [R, SW]
  const float* attn_old; const float* cov_old; float* cov_new;   // [R, T]
  const float* x_b; const float* dec_b; const float* att_b;
  float* dfeat;                                      // [R, 512]
  float* dpack;                                      // [B, K/2, 512, 2] (optional): the layout attn_fused.cu copies
  int R, K, T, V, use_cov;
};

namespace {

__global__ void __cluster_dims__(DC_CL, 1, 1) __launch_bounds__(DC_NT, 1)
dec_cell_kernel(const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2,
                const __grid_constant__ CUtensorMap tmW3, DcArgs a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* ring = smem + OFF_RING;
  float* A1 = reinterpret_cast<float*>(smem + OFF_A1);
  float* A2x = reinterpret_cast<float*>(smem + OFF_A2X);
  float* A2h = reinterpret_cast<float*>(smem + OFF_A2H);
  float* A3 = reinterpret_cast<float*>(smem + OFF_A3);
  float* red = reinterpret_cast<float*>(smem + OFF_RED);
  float* cprev = reinterpret_cast<float*>(smem + OFF_CPREV);
  float* xout = reinterpret_cast<float*>(smem + OFF_XOUT);
  float* chout = reinterpret_cast<float*>(smem + OFF_CHOUT);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + OFF_BAR);  // [4] weight stage landed
  uint64_t* empty = full + DC_STAGES;                            // [4] ... consumed by the 8 compute warps
  uint64_t* in_bar = empty + DC_STAGES;                          // the block's inputs landed
  uint64_t* x_bar = in_bar + 1;                                  // x of all 8 CTAs landed
  uint64_t* ch_bar = x_bar + 1;                                  // (c, h) of all 8 CTAs landed
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint32_t c;                                                    // rank in the cluster = which column slices
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(c));
  const int r0 = (int)(blockIdx.x / DC_CL) * DC_RB;

  // a block whose articles have all finished has nothing to do (the same decision in all 8 CTAs)
  int* rowok = reinterpret_cast<int*>(smem + OFF_ROWOK);
  int mine = 0;
  if (tid < DC_RB) {
    const int r = r0 + tid;
    mine = (r < a.R && a.active[r / a.K]) ? 1 : 0;
    rowok[tid] = mine;
  }
  if (!__syncthreads_or(mine)) return;

  if (tid == 0) {
    for (int s = 0; s < DC_STAGES; ++s) { dc_init(&full[s], 1); dc_init(&empty[s], DC_NC / 32); }
    dc_init(in_bar, 1); dc_init(x_bar, 1); dc_init(ch_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    dc_expect(x_bar, DC_CL * DC_RB * DC_N1 * 4);                 // 8 x 1 KB
    dc_expect(ch_bar, DC_CL * DC_RB * 64 * 4);                   // 8 x 4 KB
  }
  __syncthreads();
  dc_cluster_sync();                                             // every CTA's barriers exist before anybody signals them

  if (warp == DC_NC / 32) {
    // ===== producer warp: the block's inputs, then the weight stream =====
    if (lane < DC_RB) {
      const int r = min(r0 + lane, a.R - 1);
      const int art = r / a.K, rp = art * a.K + a.sel_parent[r];
      int tok = a.sel_token[r];
      if (tok >= a.V) tok = 0;                                   // beam_search.py:112-113  OOV -> [UNK]
      const float* so = a.state_old + (size_t)rp * DC_SW;
      if (lane == 0) dc_expect(in_bar, DC_RB * (DC_E + DC_A + DC_H + 32) * 4);
      __syncwarp(0x0000ffffu);
      dc_bulk(A1 + lane * DC_K1, a.emb + (size_t)tok * DC_E, DC_E * 4, in_bar);           // model.py:214
      dc_bulk(A1 + lane * DC_K1 + DC_E, so, DC_A * 4, in_bar);                             // previous context
      dc_bulk(A2h + lane * DC_H, so + DC_A + DC_H, DC_H * 4, in_bar);
      dc_bulk(cprev + lane * 32, so + DC_A + 32 * c, 32 * 4, in_bar);
    }
    if (lane == 0) {
      uint32_t n = 0;
      auto stage = [&](uint32_t bytes) -> unsigned char* {
        const uint32_t s = n % DC_STAGES;
        dc_wait(&empty[s], ((n / DC_STAGES) & 1) ^ 1);
        dc_expect(&full[s], bytes);
        return ring + (size_t)s * DC_STAGE;
      };
      for (int q = 0; q < DC_CH1; ++q, ++n) {                    // x: 128 k per chunk, 4 boxes of 16 columns
        unsigned char* dst = stage(4 * DC_N1 * 128);
        for (int b = 0; b < 4; ++b) dc_tma(dst + b * DC_N1 * 128, &tmW1, 128 * q + 32 * b, DC_N1 * c, &full[n % DC_STAGES]);
      }
      for (int i = 0; i < DC_CH2; ++i, ++n) {                    // gates: 32 k per chunk; the h part first (x arrives later)
        const int q = (i + 4) % DC_CH2;
        unsigned char* dst = stage(DC_N2 * 128);
        dc_tma(dst, &tmW2, 32 * q, DC_N2 * c, &full[n % DC_STAGES]);
      }
      for (int q = 0; q < DC_CH3; ++q, ++n) {                    // attention features: 64 k per chunk, 2 boxes
        unsigned char* dst = stage(2 * DC_N3 * 128);
        for (int b = 0; b < 2; ++b) dc_tma(dst + b * DC_N3 * 128, &tmW3, 64 * q + 32 * b, DC_N3 * c, &full[n % DC_STAGES]);
      }
    }
    dc_cluster_sync();                                           // (matches the compute warps' final cluster barrier)
    return;
  }

  if (warp == DC_NC / 32 + 1) {
    // ===== one warp: coverage of the parent + its attention (:114), rows 2c and 2c + 1 of the block =====
    if (a.use_cov) {
      for (int i = 0; i < 2; ++i) {
        const int row = 2 * (int)c + i, r = r0 + row;
        if (!rowok[row]) continue;
        const int art = r / a.K;
        const int rp = art * a.K + a.sel_parent[r], len = a.lens[art];
        const float* co = a.cov_old + (size_t)rp * a.T;
        const float* ao = a.attn_old + (size_t)rp * a.T;
        float* cn = a.cov_new + (size_t)r * a.T;
        if ((a.T & 3) == 0) {
          for (int t4 = lane; t4 < (len + 3) / 4; t4 += 128) {        // four float4 pairs in flight per lane
            float4 x[4], y[4];
#pragma unroll
            for (int q = 0; q < 4; ++q)
              if (t4 + 32 * q < (len + 3) / 4) { x[q] = reinterpret_cast<const float4*>(co)[t4 + 32 * q]; y[q] = reinterpret_cast<const float4*>(ao)[t4 + 32 * q]; }
#pragma unroll
            for (int q = 0; q < 4; ++q)
              if (t4 + 32 * q < (len + 3) / 4)
                reinterpret_cast<float4*>(cn)[t4 + 32 * q] = make_float4(x[q].x + y[q].x, x[q].y + y[q].y, x[q].z + y[q].z, x[q].w + y[q].w);
          }
        } else {
          for (int t = lane; t < len; t += 32) cn[t] = co[t] + ao[t];
        }
      }
    }
    dc_cluster_sync();                                           // (matches the compute warps' final cluster barrier)
    return;
  }

  // ===== compute warps =====
  // biases of this thread's outputs, requested now: the epilogues must not wait for an L2 round trip
  const float bias1 = a.x_b[DC_N1 * c + (tid & 15)];
  float bias2[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) bias2[g] = a.dec_b[g * DC_H + 32 * c + (tid & 31)];
  const float bias3 = a.att_b[DC_N3 * c + (tid & 63)];
  uint32_t n = 0;                                                // weight chunk counter (same sequence as the producer)
  auto chunk_wait = [&]() -> const unsigned char* {
    const uint32_t s = n % DC_STAGES;
    dc_wait(&full[s], (n / DC_STAGES) & 1);
    return ring + (size_t)s * DC_STAGE;
  };
  auto chunk_done = [&]() {
    __syncwarp();
    if (lane == 0) dc_arrive(&empty[n % DC_STAGES]);
    ++n;
  };
  dc_wait(in_bar, 0);

  // ---------------- stage 1: x = [emb; ctx] . M + b, columns [16c, +16) ----------------
  {
    const int kg = warp, rh = lane >> 4;                         // 16 k groups; lane = (column, row half)
    const int col[1] = {lane & 15};
    unsigned long long acc[1][8];
#pragma unroll
    for (int r = 0; r < 8; ++r) acc[0][r] = 0ull;
    for (int q = 0; q < DC_CH1; ++q) {
      const unsigned char* w = chunk_wait() + (kg >> 2) * (DC_N1 * 128);     // the box of this warp's 8 k
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int jj = 2 * (kg & 3) + i;
        dc_step<1>(acc, A1 + (rh * 8) * DC_K1 + 128 * q + 32 * (kg >> 2) + 4 * jj, DC_K1, w, col, jj);
      }
      chunk_done();
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) red[(kg * DC_RB + rh * 8 + r) * DC_N1 + col[0]] = dc_sum2(acc[0][r]);
    dc_sync_compute();
    if (tid < DC_RB * DC_N1) {                                   // 256 outputs; k groups in fixed order
      const int row = tid >> 4, cc = tid & 15;
      float v = 0.f;
#pragma unroll
      for (int g = 0; g < 16; ++g) v += red[(g * DC_RB + row) * DC_N1 + cc];
      v += bias1;
      xout[row * DC_N1 + cc] = v;
      const int r = r0 + row;
      if (rowok[row]) a.state_new[(size_t)r * DC_SW + DC_A + 2 * DC_H + DC_N1 * c + cc] = v;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // ordinary stores -> bulk-copy reads
    dc_sync_compute();
    if (tid < DC_CL) {                                           // this CTA's 16 x 16 block of x to every CTA of the cluster
      const uint32_t dst = dc_mapa(dc_u32(A2x + c * (DC_RB * DC_N1)), tid), rbar = dc_mapa(dc_u32(x_bar), tid);
      asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst), "r"(dc_u32(xout)), "n"(DC_RB * DC_N1 * 4), "r"(rbar) : "memory");
    }
  }

  // ---------------- stage 2: gates of units [32c, +32), LSTM cell ----------------
  {
    const int kg = warp & 3, rh = (warp >> 2) & 1, ch = warp >> 3;    // 4 k groups x 2 row halves x 2 column halves
    const int col[2] = {64 * ch + lane, 64 * ch + lane + 32};    // local column = gate * 32 + unit
    unsigned long long acc[2][8];
#pragma unroll
    for (int cc = 0; cc < 2; ++cc)
#pragma unroll
      for (int r = 0; r < 8; ++r) acc[cc][r] = 0ull;
    for (int i = 0; i < DC_CH2; ++i) {
      const int q = (i + 4) % DC_CH2;                            // k = 32 q: q >= 4 is the h part, q < 4 the x part
      if (i == DC_CH2 - 4) dc_wait(x_bar, 0);                    // x of all 8 CTAs (sent while the h part was multiplied)
      const unsigned char* w = chunk_wait();
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int jj = 2 * kg + u, k = 32 * q + 4 * jj;
        if (q >= 4) dc_step<2>(acc, A2h + (rh * 8) * DC_H + (k - DC_E), DC_H, w, col, jj);
        else dc_step<2>(acc, A2x + (k >> 4) * (DC_RB * DC_N1) + (rh * 8) * DC_N1 + (k & 15), DC_N1, w, col, jj);
      }
      chunk_done();
    }
#pragma unroll
    for (int cc = 0; cc < 2; ++cc)
#pragma unroll
      for (int r = 0; r < 8; ++r) red[(kg * DC_RB + rh * 8 + r) * DC_N2 + col[cc]] = dc_sum2(acc[cc][r]);
    dc_sync_compute();
    {                                                            // (row, unit): k groups in fixed order, then the cell
      const int row = tid >> 5, u = tid & 31;
      float z[4];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float v = 0.f;
#pragma unroll
        for (int q = 0; q < 4; ++q) v += red[(q * DC_RB + row) * DC_N2 + g * 32 + u];
        z[g] = v + bias2[g];
      }
      // LSTMCell: gates i, j, f, o; forget bias 1.0 (model.py:138)
      const float cn = dc_sigmoid(z[2] + 1.0f) * cprev[row * 32 + u] + dc_sigmoid(z[0]) * tanhf(z[1]);
      const float hn = dc_sigmoid(z[3]) * tanhf(cn);
      chout[row * 64 + u] = cn;
      chout[row * 64 + 32 + u] = hn;
      const int r = r0 + row;
      if (rowok[row]) {
        a.state_new[(size_t)r * DC_SW + DC_A + 32 * c + u] = cn;
        a.state_new[(size_t)r * DC_SW + DC_A + DC_H + 32 * c + u] = hn;
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    dc_sync_compute();
    if (tid < DC_CL) {                                           // this CTA's 16 x (32 c | 32 h) to every CTA of the cluster
      const uint32_t dst = dc_mapa(dc_u32(A3 + c * (DC_RB * 64)), tid), rbar = dc_mapa(dc_u32(ch_bar), tid);
      asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst), "r"(dc_u32(chout)), "n"(DC_RB * 64 * 4), "r"(rbar) : "memory");
    }
  }

  // ---------------- stage 3: attention features d = [c; h] . W_s + b, columns [64c, +64) ----------------
  {
    const int kg = warp & 3, rh = (warp >> 2) & 1, ch = warp >> 3;
    const int col[1] = {32 * ch + lane};
    unsigned long long acc[1][8];
#pragma unroll
    for (int r = 0; r < 8; ++r) acc[0][r] = 0ull;
    dc_wait(ch_bar, 0);                                          // (c, h) of all 8 CTAs
    for (int q = 0; q < DC_CH3; ++q) {                           // chunk q = the 64 values that came from CTA q
      const unsigned char* w = chunk_wait() + (kg >> 1) * (DC_N3 * 128);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int jj = 4 * (kg & 1) + u;
        dc_step<1>(acc, A3 + q * (DC_RB * 64) + (rh * 8) * 64 + 32 * (kg >> 1) + 4 * jj, 64, w, col, jj);
      }
      chunk_done();
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) red[(kg * DC_RB + rh * 8 + r) * DC_N3 + col[0]] = dc_sum2(acc[0][r]);
    dc_sync_compute();
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int o = tid + it * DC_NC, row = o >> 6, cc = o & 63;
      float v = 0.f;
#pragma unroll
      for (int q = 0; q < 4; ++q) v += red[(q * DC_RB + row) * DC_N3 + cc];
      v += bias3;
      const int r = r0 + row;
      if (rowok[row]) {
        const int f = DC_N3 * c + cc;
        a.dfeat[(size_t)r * DC_A + f] = v;
        if (a.dpack) {                                           // [article][row pair][feature][2], scaled: attn_fused.cu
          const int art = r / a.K, k = r - art * a.K;
          a.dpack[(((size_t)art * (a.K / 2) + (k >> 1)) * DC_A + f) * 2 + (k & 1)] = kS * v;
        }
      }
    }
  }
  dc_cluster_sync();                                             // nobody leaves while its shared memory can still be a copy's target
}

// ---------------------------------------------------------------------------------------------
// output = [h; ctx] . AttnOutputProjection + b (:173) and p_gen (:166-167), after the attention pass.
// CTA (cb, rb): columns [32 cb, +32) of the 16 rows of block rb; the CTAs with cb == 0 also compute p_gen.
// ---------------------------------------------------------------------------------------------
constexpr int DO_K = DC_H + DC_A, DO_N = 32, DO_CH = DO_K / 128;           // 768, 32 columns per CTA, 6 chunks
constexpr int DO_OFF_A = DC_STAGES * DC_STAGE;                              // [16][768]  h | ctx
constexpr int DO_OFF_RED = DO_OFF_A + DC_RB * DO_K * 4;                     // [8 k groups][16][32]
constexpr int DO_OFF_BAR = DO_OFF_RED + 8 * DC_RB * DO_N * 4;
constexpr int DO_OFF_ROWOK = DO_OFF_BAR + 128;
constexpr int DO_SMEM = DO_OFF_ROWOK + 64;
constexpr int DO_NT = DC_NC + 32;                                           // compute threads + one producer warp

struct DoArgs {
  const int* active;
  const float* state;                                // [R, SW] of this step: ctx | c | h | x
  const float* out_b; const float* pg_w; const float* pg_b;
  float* out; float* out_tf32;                       // [R, 256]; out_tf32 optional (operand of the tensor-core pass)
  float* pgen;                                       // [R] (nullptr: pointer_gen off)
  int R, K;
};

__global__ void __launch_bounds__(DO_NT, 1)
dec_out_kernel(const __grid_constant__ CUtensorMap tmW, DoArgs a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* ring = smem;
  float* A = reinterpret_cast<float*>(smem + DO_OFF_A);
  float* red = reinterpret_cast<float*>(smem + DO_OFF_RED);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + DO_OFF_BAR);
  uint64_t* empty = full + DC_STAGES;
  uint64_t* in_bar = empty + DC_STAGES;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int cb = blockIdx.x, r0 = blockIdx.y * DC_RB;
  int* rowok = reinterpret_cast<int*>(smem + DO_OFF_ROWOK);
  int mine = 0;
  if (tid < DC_RB) {
    const int r = r0 + tid;
    mine = (r < a.R && a.active[r / a.K]) ? 1 : 0;
    rowok[tid] = mine;
  }
  if (tid == 0) {
    for (int s = 0; s < DC_STAGES; ++s) { dc_init(&full[s], 1); dc_init(&empty[s], DC_NC / 32); }
    dc_init(in_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (!__syncthreads_or(mine)) return;
  if (warp == DC_NC / 32) {
    if (lane < DC_RB) {
      const int r = min(r0 + lane, a.R - 1);
      const float* s = a.state + (size_t)r * DC_SW;
      if (lane == 0) dc_expect(in_bar, DC_RB * DO_K * 4);
      __syncwarp(0x0000ffffu);
      dc_bulk(A + lane * DO_K, s + DC_A + DC_H, DC_H * 4, in_bar);           // cell output h
      dc_bulk(A + lane * DO_K + DC_H, s, DC_A * 4, in_bar);                  // context
    }
    if (lane == 0) {
      for (uint32_t n = 0; n < (uint32_t)DO_CH; ++n) {
        const uint32_t s = n % DC_STAGES;
        dc_wait(&empty[s], ((n / DC_STAGES) & 1) ^ 1);
        dc_expect(&full[s], 4 * DO_N * 128);
        for (int b = 0; b < 4; ++b) dc_tma(ring + (size_t)s * DC_STAGE + b * DO_N * 128, &tmW, 128 * n + 32 * b, DO_N * cb, &full[s]);
      }
    }
    return;
  }
  const int kg = warp & 7, rh = warp >> 3;                       // 8 k groups x 2 row halves; lane = column
  const int col[1] = {lane};
  const float bias = a.out_b[DO_N * cb + lane];
  unsigned long long acc[1][8];
#pragma unroll
  for (int r = 0; r < 8; ++r) acc[0][r] = 0ull;
  dc_wait(in_bar, 0);
  for (uint32_t n = 0; n < (uint32_t)DO_CH; ++n) {
    const uint32_t s = n % DC_STAGES;
    dc_wait(&full[s], (n / DC_STAGES) & 1);
    const unsigned char* w = ring + (size_t)s * DC_STAGE + (kg >> 1) * (DO_N * 128);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int jj = 4 * (kg & 1) + u;
      dc_step<1>(acc, A + (rh * 8) * DO_K + 128 * n + 32 * (kg >> 1) + 4 * jj, DO_K, w, col, jj);
    }
    __syncwarp();
    if (lane == 0) dc_arrive(&empty[s]);
  }
#pragma unroll
  for (int r = 0; r < 8; ++r) red[(kg * DC_RB + rh * 8 + r) * DO_N + lane] = dc_sum2(acc[0][r]);
  dc_sync_compute();
  {
    const int row = tid >> 5, cc = tid & 31, r = r0 + row;       // 512 outputs, one per thread; k groups in fixed order
    float v = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) v += red[(q * DC_RB + row) * DO_N + cc];
    v += bias;
    if (rowok[row]) {
      a.out[(size_t)r * DC_H + DO_N * cb + cc] = v;
      if (a.out_tf32) {
        uint32_t t;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(v));
        a.out_tf32[(size_t)r * DC_H + DO_N * cb + cc] = __uint_as_float(t);
      }
    }
  }
  if (cb == 0 && a.pgen) {                                       // p_gen: warp w takes row w
    {
      const int r = r0 + warp;
      if (!rowok[warp]) return;
      const float4* s4 = reinterpret_cast<const float4*>(a.state + (size_t)r * DC_SW);
      const float4* w4 = reinterpret_cast<const float4*>(a.pg_w);
      float4 sv[DC_SW / 128], wv[DC_SW / 128];
#pragma unroll
      for (int q = 0; q < DC_SW / 128; ++q) { sv[q] = s4[lane + 32 * q]; wv[q] = __ldg(w4 + lane + 32 * q); }
      float p = 0.f;
#pragma unroll
      for (int q = 0; q < DC_SW / 128; ++q) {
        p = fmaf(sv[q].x, wv[q].x, p); p = fmaf(sv[q].y, wv[q].y, p);
        p = fmaf(sv[q].z, wv[q].z, p); p = fmaf(sv[q].w, wv[q].w, p);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) p += __shfl_xor_sync(0xffffffffu, p, o);
      if (lane == 0) a.pgen[r] = dc_sigmoid(p + a.pg_b[0]);
    }
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
namespace {

// gate rows of lstm_cell/kernel^T regrouped so that CTA c's 128 columns (4 gates x units [32c, +32)) are contiguous
__global__ void dc_permute_gates_kernel(const float* __restrict__ in, float* __restrict__ out) {
  const int row = blockIdx.x;                                    // new row = c * 128 + g * 32 + u
  const int c = row >> 7, g = (row >> 5) & 3, u = row & 31;
  const float* src = in + (size_t)(g * DC_H + 32 * c + u) * DC_K2;
  for (int k = threadIdx.x; k < DC_K2; k += blockDim.x) out[(size_t)row * DC_K2 + k] = src[k];
}
// k of Attention/Linear/Matrix^T regrouped into the order the (c, h) exchange delivers: k' = 64 s + j is the j-th of
// the 32 c units, then the 32 h units, that CTA s computed
__global__ void dc_permute_state_kernel(const float* __restrict__ in, float* __restrict__ out) {
  const int row = blockIdx.x;
  for (int kp = threadIdx.x; kp < DC_K3; kp += blockDim.x) {
    const int s = kp >> 6, j = kp & 63;
    const int k = j < 32 ? 32 * s + j : DC_H + 32 * s + (j - 32);
    out[(size_t)row * DC_K3 + kp] = in[(size_t)row * DC_K3 + k];
  }
}

typedef CUresult (*DcEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                               const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                               CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

std::string dc_make_map(DcEncodeFn fn, CUtensorMap* map, const float* base, int rows, int K, int box_rows) {
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)K * sizeof(float)};
  cuuint32_t box[2] = {32, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r);
  return "";
}

}  // namespace

std::string DecCell::setup(const float* x_MT, const float* dec_kT, const float* att_MT, const float* out_MT, float* k2,
                           float* k3, cudaStream_t st) {
  dc_permute_gates_kernel<<<4 * DC_H, 128, 0, st>>>(dec_kT, k2);
  dc_permute_state_kernel<<<DC_A, 128, 0, st>>>(att_MT, k3);
  if (cudaGetLastError() != cudaSuccess) return "permuting the decoder weights failed";
  void* fnp = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fnp, cudaEnableDefault, &q) != cudaSuccess || !fnp)
    return "cuTensorMapEncodeTiled is not available from the driver";
  DcEncodeFn fn = (DcEncodeFn)fnp;
  std::string err = dc_make_map(fn, &tmW1, x_MT, DC_E, DC_K1, DC_N1);
  if (err.empty()) err = dc_make_map(fn, &tmW2, k2, 4 * DC_H, DC_K2, DC_N2);
  if (err.empty()) err = dc_make_map(fn, &tmW3, k3, DC_A, DC_K3, DC_N3);
  if (err.empty()) err = dc_make_map(fn, &tmWo, out_MT, DC_H, DO_K, DO_N);
  if (!err.empty()) return err;
  if (cudaFuncSetAttribute(dec_cell_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DC_SMEM) != cudaSuccess ||
      cudaFuncSetAttribute(dec_out_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DO_SMEM) != cudaSuccess) {
    cudaGetLastError();
    return "cudaFuncSetAttribute(dec_cell_kernel) failed";
  }
  ready = true;
  return "";
}

int DecCell::launch_cell(const Dims& d, int B, const StepBufs& b, const float* x_b, const float* dec_b, const float* att_b,
                         float* dfeat, float* dpack, cudaStream_t st) const {
  DcArgs a;
  a.lens = b.lens; a.active = b.active; a.sel_parent = b.sel_parent; a.sel_token = b.sel_token; a.emb = b.emb;
  a.state_old = b.state_old; a.state_new = b.state_new; a.attn_old = b.attn_old; a.cov_old = b.cov_old; a.cov_new = b.cov_new;
  a.x_b = x_b; a.dec_b = dec_b; a.att_b = att_b; a.dfeat = dfeat; a.dpack = dpack;
  a.R = B * d.K; a.K = d.K; a.T = d.T; a.V = d.V; a.use_cov = d.use_cov;
  const int blocks = (a.R + DC_RB - 1) / DC_RB;
  dec_cell_kernel<<<blocks * DC_CL, DC_NT, DC_SMEM, st>>>(tmW1, tmW2, tmW3, a);
  return 1;
}

int DecCell::launch_out(const Dims& d, int B, const int* active, const float* state, const float* out_b, const float* pg_w,
                        const float* pg_b, float* out, float* out_tf32, float* pgen, cudaStream_t st) const {
  DoArgs a;
  a.active = active; a.state = state; a.out_b = out_b; a.pg_w = pg_w; a.pg_b = pg_b; a.out = out; a.out_tf32 = out_tf32;
  a.pgen = d.pointer_gen ? pgen : nullptr; a.R = B * d.K; a.K = d.K;
  dim3 grid(DC_H / DO_N, (a.R + DC_RB - 1) / DC_RB);
  dec_out_kernel<<<grid, DO_NT, DO_SMEM, st>>>(tmWo, a);
  return 1;
}

}  // namespace pgn
