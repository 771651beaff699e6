// One attention pass of the decode step in ONE launch (attention_decoder.py:79-129): energies with coverage,
// masked softmax, re-normalisation, context vector -- for every beam row of every live article.
//
//   e[r,t]   = sum_a v_a tanh(F[b,t,a] + d[r,a] + cov[r,t] w_c,a)             (:105-117)
//   attn[r,:] = softmax(e[r,:len]) re-normalised                               (:96-101)
//   ctx[r,:]  = sum_t attn[r,t] h[b,t,:]                                       (:126-127)
//
// The encoder features F and the encoder states h of an article (2 x T x 2 KB) are each read from HBM exactly once
// per step and shared by the article's K beam rows.  Work item = (article, segment of 32 encoder positions); the
// segmentation depends on the article alone, so an article's arithmetic is the same whatever shares the batch.  A
// persistent grid of one CTA per SM walks the items; every CTA streams the F tile and the h tile of its items through
// a ring of three 64.5 KB shared-memory buffers filled by cp.async.bulk row copies (rows padded by 16 B so that
// "lane = position" reads are conflict-free) completing on mbarriers, so that the tiles of the next item are in
// flight while this one is computed:
//
//   E phase  warp = (pair of beam rows, slice of the 512 features), lane = position: packed FFMA2 over the row pair,
//            tanh(x) = 1 - 2 / (1 + 2^(2 log2e x)) on the special-function unit, four features share one reciprocal
//            (1.25 MUFU operations per tanh); the slices' partial sums are added in fixed order;
//   softmax  inside the segment (its own maximum), p = exp(e - m_seg);
//   C phase  thread = feature: ctx_seg[k][a] = sum_t p[t][k] h[t][a].
//
// Every segment writes (m_seg, sum p, ctx_seg) to a small scratch; the CTA that completes an article's LAST segment
// (a counter per article, self-resetting) merges the segments in segment order -- M = max m_seg, S = sum s_seg
// e^(m_seg - M) -- and writes attn (softmax, then the reference's re-normalisation) and the context vector.
#include <cuda_runtime.h>

#include <cfloat>
#include <cstdint>

#include "kernels.h"

namespace pgn {

namespace {

constexpr int AF_A = 512, AF_SEG = 32, AF_NT = 512, AF_NW = AF_NT / 32;
constexpr int AF_ROWF = AF_A + 4;                    // row pitch in floats: 2 KB + 16 B
constexpr int AF_TILE = AF_SEG * AF_ROWF * 4;        // bytes of one ring buffer
constexpr float kS = 2.885390081777927f;             // 2 log2(e): e^{2x} = 2^{kS x}

template <int K> struct AfCfg {
  static constexpr int NPAIR = K / 2;
  static constexpr int NSLICE = AF_NW / NPAIR;       // warps (feature slices) per row pair
  static constexpr int SLICE = AF_A / NSLICE;        // features per warp
  static constexpr int OFF_PART = 3 * AF_TILE;                       // [NSLICE][K][32] partial energies
  static constexpr int OFF_DS = OFF_PART + NSLICE * K * 32 * 4;      // [NPAIR][512][2] kS * decoder features
  static constexpr int OFF_VS = OFF_DS + AF_A * K * 4;               // [512] -2 v
  static constexpr int OFF_WS = OFF_VS + AF_A * 4;                   // [512] kS w_c
  static constexpr int OFF_P = OFF_WS + AF_A * 4;                    // [32][K] p = exp(e - m_seg)
  static constexpr int OFF_RED = OFF_P + 32 * K * 4;                 // [AF_NW][K] + [3][K] merge scratch
  static constexpr int OFF_BAR = OFF_RED + (AF_NW * K + 4 * K) * 4;  // 3 mbarriers + flag
  static constexpr int SMEM = OFF_BAR + 64;
};

__device__ __forceinline__ uint32_t af_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned long long af_pk2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void af_upk2(unsigned long long v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ unsigned long long af_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ unsigned long long af_mul2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ unsigned long long af_add2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float af_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float af_rcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ void af_bar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = af_u32(bar);
  uint32_t ok = 0, spins = 0;
  while (!ok) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    if (!ok && ++spins > (1u << 26)) __trap();       // a protocol bug becomes an error, not a hang
  }
}

struct AfArgs {
  const int* lens; const int* active;
  const float* enc_feat; const float* enc_states;    // [B, T, 512]
  const float* dfeat;                                // [B*K, 512]
  const float* cov;                                  // [B*K, T] (nullptr: no coverage term)
  const float* v; const float* w_c;                  // [512]
  float* e;                                          // [B*K, T] energies (scratch)
  float* pctx;                                       // [B, NSEG, K, 512] per-segment context partials (scratch)
  float* segstat;                                    // [B, NSEG, K, 2] per-segment (max, sum) (scratch)
  int* counters;                                     // [B] segments completed; zero before the launch, zero after it
  float* attn;                                       // [B*K, T]
  float* ctx; int ctx_ld;                            // context vector of row r at ctx + r * ctx_ld
  int B, T, nseg_max;
};

// Warp 0 issues the row copies of one tile (rows [t0, t0 + n) of `src`, 2 KB each) into a ring buffer.
__device__ __forceinline__ void af_issue_tile(float* buf, uint64_t* bar, const float* src, int n, int lane) {
  if (lane == 0) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // the buffer's last readers used ordinary loads
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(af_u32(bar)), "r"((uint32_t)n * AF_A * 4u) : "memory");
  }
  __syncwarp();
  if (lane < n)
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(af_u32(buf + lane * AF_ROWF)), "l"(src + (size_t)lane * AF_A), "n"(AF_A * 4), "r"(af_u32(bar)) : "memory");
}

template <int K>
__global__ void __launch_bounds__(AF_NT, 1)
attn_fused_kernel(AfArgs a) {
  using C = AfCfg<K>;
  extern __shared__ __align__(128) unsigned char smraw[];
  float* tiles = reinterpret_cast<float*>(smraw);
  float* part = reinterpret_cast<float*>(smraw + C::OFF_PART);
  float* ds = reinterpret_cast<float*>(smraw + C::OFF_DS);
  float* vs = reinterpret_cast<float*>(smraw + C::OFF_VS);
  float* ws = reinterpret_cast<float*>(smraw + C::OFF_WS);
  float* psm = reinterpret_cast<float*>(smraw + C::OFF_P);
  float* red = reinterpret_cast<float*>(smraw + C::OFF_RED);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smraw + C::OFF_BAR);
  int* flag = reinterpret_cast<int*>(bars + 3);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int T = a.T, NSEG = a.nseg_max;
  const int n_items = a.B * NSEG;

  // the items of this CTA: i = blockIdx.x, + gridDim.x, ... restricted to live articles' existing segments
  auto valid = [&](int i) -> bool {
    const int art = i / NSEG, seg = i - art * NSEG;
    return a.active[art] && seg * AF_SEG < a.lens[art];
  };
  auto next_item = [&](int i) -> int {
    i += gridDim.x;
    while (i < n_items && !valid(i)) i += gridDim.x;
    return i;
  };
  int it0 = (int)blockIdx.x;
  while (it0 < n_items && !valid(it0)) it0 += gridDim.x;
  if (it0 >= n_items) return;
  int it1 = next_item(it0);

  if (tid == 0) {
    for (int b = 0; b < 3; ++b) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(af_u32(&bars[b])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = tid; i < AF_A; i += AF_NT) { vs[i] = -2.0f * a.v[i]; ws[i] = a.cov ? kS * a.w_c[i] : 0.f; }
  __syncthreads();
  float vsum = 0.f;                                  // sum_a v_a: e = sum v - 2 sum v / (1 + e^{2x})
  for (int i = lane; i < AF_A; i += 32) vsum += vs[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) vsum += __shfl_xor_sync(0xffffffffu, vsum, o);
  vsum *= -0.5f;

  auto item_geom = [&](int i, int& art, int& seg, int& n) {
    art = i / NSEG; seg = i - art * NSEG;
    n = min(AF_SEG, a.lens[art] - seg * AF_SEG);
  };
  // tile loads are numbered: load 2j = F tile of the CTA's j-th item, 2j + 1 = its h tile; load n uses buffer n % 3
  uint32_t n_load = 0;                               // next load to issue (warp 0)
  auto issue = [&](int i, bool states) {
    int art, seg, n;
    item_geom(i, art, seg, n);
    const float* src = (states ? a.enc_states : a.enc_feat) + ((size_t)art * T + (size_t)seg * AF_SEG) * AF_A;
    const uint32_t b = n_load % 3;
    af_issue_tile(tiles + (size_t)b * (AF_TILE / 4), &bars[b], src, n, lane);
    ++n_load;
  };
  if (warp == 0) {
    issue(it0, false);
    issue(it0, true);
    if (it1 < n_items) issue(it1, false);
  }
  // per-warp roles
  const int pair = warp % C::NPAIR, slice = warp / C::NPAIR;
  // decoder features and coverage of the first item
  float dreg[K];
  float cv0 = 0.f, cv1 = 0.f;
  {
    int art, seg, n;
    item_geom(it0, art, seg, n);
#pragma unroll
    for (int k = 0; k < K; ++k) dreg[k] = a.dfeat[(size_t)(art * K + k) * AF_A + tid];
    if (a.cov && lane < n) {
      cv0 = a.cov[(size_t)(art * K + 2 * pair) * T + seg * AF_SEG + lane];
      cv1 = a.cov[(size_t)(art * K + 2 * pair + 1) * T + seg * AF_SEG + lane];
    }
  }
  const unsigned long long kS2 = af_pk2(kS, kS), one2 = af_pk2(1.0f, 1.0f);
  uint32_t j = 0;                                    // index of the current item in this CTA's sequence
  for (int it = it0; it < n_items; it = it1, it1 = next_item(it1), ++j) {
    int art, seg, n;
    item_geom(it, art, seg, n);
    const int t0 = seg * AF_SEG;
    // ds of this item (the E phase of the previous item is over: see the barrier at the end of the loop body)
#pragma unroll
    for (int p = 0; p < C::NPAIR; ++p)
      *reinterpret_cast<float2*>(ds + ((size_t)p * AF_A + tid) * 2) = make_float2(kS * dreg[2 * p], kS * dreg[2 * p + 1]);
    const unsigned long long cv2 = af_pk2(cv0, cv1);
    // prefetch the next item's decoder features and coverage into registers
    if (it1 < n_items) {
      int art1, seg1, n1;
      item_geom(it1, art1, seg1, n1);
#pragma unroll
      for (int k = 0; k < K; ++k) dreg[k] = a.dfeat[(size_t)(art1 * K + k) * AF_A + tid];
      cv0 = cv1 = 0.f;
      if (a.cov && lane < n1) {
        cv0 = a.cov[(size_t)(art1 * K + 2 * pair) * T + seg1 * AF_SEG + lane];
        cv1 = a.cov[(size_t)(art1 * K + 2 * pair + 1) * T + seg1 * AF_SEG + lane];
      }
    }
    __syncthreads();                                 // ds complete
    // ---------------- E phase ----------------
    {
      const uint32_t ld = 2 * j, b = ld % 3;
      af_bar_wait(&bars[b], (ld / 3) & 1);
      const float* Frow = tiles + (size_t)b * (AF_TILE / 4) + lane * AF_ROWF + slice * C::SLICE;
      const float4* vs4 = reinterpret_cast<const float4*>(vs + slice * C::SLICE);
      const float4* ws4 = reinterpret_cast<const float4*>(ws + slice * C::SLICE);
      const float4* ds4 = reinterpret_cast<const float4*>(ds + ((size_t)pair * AF_A + slice * C::SLICE) * 2);
      unsigned long long acc2 = 0ull;
#pragma unroll 2
      for (int c = 0; c < C::SLICE / 4; ++c) {
        const float4 F = *reinterpret_cast<const float4*>(Frow + 4 * c);
        const float4 vv = vs4[c], ww = ws4[c];
        const float4 d01 = ds4[2 * c], d23 = ds4[2 * c + 1];
        const float fv[4] = {F.x, F.y, F.z, F.w}, wv[4] = {ww.x, ww.y, ww.z, ww.w};
        const unsigned long long dd[4] = {af_pk2(d01.x, d01.y), af_pk2(d01.z, d01.w), af_pk2(d23.x, d23.y), af_pk2(d23.z, d23.w)};
        unsigned long long Aq[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const unsigned long long y = af_fma2(cv2, af_pk2(wv[i], wv[i]), af_fma2(af_pk2(fv[i], fv[i]), kS2, dd[i]));
          float y0, y1;
          af_upk2(y, y0, y1);
          // x is clamped at 10 (tanh(10) rounds to 1.0f): 1 + e^{2x} <= 1 + e^20 and the product of four stays below 6e34
          Aq[i] = af_add2(one2, af_pk2(af_ex2(fminf(y0, 10.f * kS)), af_ex2(fminf(y1, 10.f * kS))));
        }
        // the four reciprocals of a row share ONE rcp: r = 1 / (A0 A1 A2 A3), 1/A0 = r A1 A2 A3, ...
        const unsigned long long p01 = af_mul2(Aq[0], Aq[1]), p23 = af_mul2(Aq[2], Aq[3]);
        float q0, q1;
        af_upk2(af_mul2(p01, p23), q0, q1);
        const unsigned long long r = af_pk2(af_rcp(q0), af_rcp(q1));
        const unsigned long long r01 = af_mul2(r, p23), r23 = af_mul2(r, p01);
        acc2 = af_fma2(af_pk2(vv.x, vv.x), af_mul2(r01, Aq[1]), acc2);
        acc2 = af_fma2(af_pk2(vv.y, vv.y), af_mul2(r01, Aq[0]), acc2);
        acc2 = af_fma2(af_pk2(vv.z, vv.z), af_mul2(r23, Aq[3]), acc2);
        acc2 = af_fma2(af_pk2(vv.w, vv.w), af_mul2(r23, Aq[2]), acc2);
      }
      float s0, s1;
      af_upk2(acc2, s0, s1);
      part[(slice * K + 2 * pair) * 32 + lane] = s0;
      part[(slice * K + 2 * pair + 1) * 32 + lane] = s1;
    }
    __syncthreads();                                 // partial energies complete; the F buffer is free
    // the freed buffer takes the h tile of the next item
    if (warp == 0 && it1 < n_items) issue(it1, true);
    // ---------------- energies, softmax inside the segment ----------------
    if (warp < K) {
      const int k = warp;
      float e = vsum;
#pragma unroll
      for (int s = 0; s < C::NSLICE; ++s) e += part[(s * K + k) * 32 + lane];       // fixed order
      const bool in = lane < n;
      float m = in ? e : -FLT_MAX;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      const float p = in ? expf(e - m) : 0.f;
      float s = p;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      psm[lane * K + k] = p;
      if (in) a.e[(size_t)(art * K + k) * T + t0 + lane] = e;
      if (lane == 0) {
        float* st = a.segstat + ((size_t)(art * NSEG + seg) * K + k) * 2;
        st[0] = m; st[1] = s;
      }
    }
    __syncthreads();                                 // p complete
    // ---------------- C phase ----------------
    {
      const uint32_t ld = 2 * j + 1, b = ld % 3;
      af_bar_wait(&bars[b], (ld / 3) & 1);
      const float* Hc = tiles + (size_t)b * (AF_TILE / 4) + tid;
      float acc[K];
#pragma unroll
      for (int k = 0; k < K; ++k) acc[k] = 0.f;
#pragma unroll 4
      for (int t = 0; t < n; ++t) {
        const float h = Hc[t * AF_ROWF];
#pragma unroll
        for (int k4 = 0; k4 < K; k4 += 4) {
          const float4 p4 = *reinterpret_cast<const float4*>(psm + t * K + k4);
          acc[k4] = fmaf(p4.x, h, acc[k4]); acc[k4 + 1] = fmaf(p4.y, h, acc[k4 + 1]);
          acc[k4 + 2] = fmaf(p4.z, h, acc[k4 + 2]); acc[k4 + 3] = fmaf(p4.w, h, acc[k4 + 3]);
        }
      }
      float* pc = a.pctx + (size_t)(art * NSEG + seg) * K * AF_A + tid;
#pragma unroll
      for (int k = 0; k < K; ++k) pc[k * AF_A] = acc[k];
    }
    // publish the segment; the CTA that completes the article's last segment merges
    __threadfence();
    __syncthreads();                                 // the h buffer and psm are free; every partial is written
    if (warp == 0) {
      // the freed buffer takes the F tile of the item after next
      const int it2 = it1 < n_items ? next_item(it1) : n_items;
      if (it2 < n_items) issue(it2, false);
    }
    const int nseg = (a.lens[art] + AF_SEG - 1) / AF_SEG;
    if (tid == 0) *flag = (atomicAdd(a.counters + art, 1) == nseg - 1) ? 1 : 0;
    __syncthreads();
    if (*flag) {
      __threadfence();
      const int len = a.lens[art];
      float* Mk = red + AF_NW * K;                   // [K] M, [K] S, [K] s2
      float* Sk = Mk + K;
      float* S2 = Sk + K;
      if (warp < K) {                                // segment statistics -> M, S of the row (segment order)
        const int k = warp;
        float m = -FLT_MAX;
        for (int s = lane; s < nseg; s += 32) m = fmaxf(m, __ldcg(a.segstat + ((size_t)(art * NSEG + s) * K + k) * 2));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        float S = 0.f;
        if (lane == 0)
          for (int s = 0; s < nseg; ++s) {
            const float* st = a.segstat + ((size_t)(art * NSEG + s) * K + k) * 2;
            S += __ldcg(st + 1) * expf(__ldcg(st) - m);
          }
        if (lane == 0) { Mk[k] = m; Sk[k] = S; }
      }
      __syncthreads();
      // softmax over the article's positions (:98), its sum for the re-normalisation (:100-101)
      float sum[K];
#pragma unroll
      for (int k = 0; k < K; ++k) sum[k] = 0.f;
      for (int t = tid; t < len; t += AF_NT) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const size_t o = (size_t)(art * K + k) * T + t;
          const float x = expf(__ldcg(a.e + o) - Mk[k]) / Sk[k];
          a.attn[o] = x;
          sum[k] += x;
        }
      }
#pragma unroll
      for (int k = 0; k < K; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum[k] += __shfl_xor_sync(0xffffffffu, sum[k], o);
        if (lane == 0) red[warp * K + k] = sum[k];
      }
      __syncthreads();
      if (tid < K) {
        float s2 = 0.f;
        for (int w = 0; w < AF_NW; ++w) s2 += red[w * K + tid];
        S2[tid] = s2;
      }
      __syncthreads();
      for (int t = tid; t < len; t += AF_NT) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const size_t o = (size_t)(art * K + k) * T + t;
          a.attn[o] = a.attn[o] / S2[k];             // written by this thread above
        }
      }
      for (int t = len + tid; t < T; t += AF_NT) {   // the masked distribution is exactly 0 past the article (:99)
#pragma unroll
        for (int k = 0; k < K; ++k) a.attn[(size_t)(art * K + k) * T + t] = 0.f;
      }
      // context = sum_seg e^(m_seg - M) ctx_seg / S, re-normalised like attn
      float cacc[K];
#pragma unroll
      for (int k = 0; k < K; ++k) cacc[k] = 0.f;
      for (int s = 0; s < nseg; ++s) {
        const float* pc = a.pctx + (size_t)(art * NSEG + s) * K * AF_A + tid;
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const float w = expf(__ldcg(a.segstat + ((size_t)(art * NSEG + s) * K + k) * 2) - Mk[k]);
          cacc[k] = fmaf(w, __ldcg(pc + k * AF_A), cacc[k]);
        }
      }
#pragma unroll
      for (int k = 0; k < K; ++k) a.ctx[(size_t)(art * K + k) * a.ctx_ld + tid] = cacc[k] / Sk[k] / S2[k];
      if (tid == 0) a.counters[art] = 0;             // ready for the next launch
      __syncthreads();                               // red / Mk are reused by a later merge of this CTA
    }
  }
}

template <int K>
static cudaError_t af_prepare() {
  return cudaFuncSetAttribute(attn_fused_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, AfCfg<K>::SMEM);
}

}  // namespace

bool attn_fused_supported(const Dims& d) { return d.A == AF_A && (d.K == 4 || d.K == 8); }

int attn_fused_nseg(const Dims& d) { return (d.T + AF_SEG - 1) / AF_SEG; }

size_t attn_fused_pctx_floats(const Dims& d, int B) { return (size_t)B * attn_fused_nseg(d) * d.K * AF_A; }
size_t attn_fused_segstat_floats(const Dims& d, int B) { return (size_t)B * attn_fused_nseg(d) * d.K * 2; }

cudaError_t attn_fused_prepare(const Dims& d) {
  if (d.K == 4) return af_prepare<4>();
  if (d.K == 8) return af_prepare<8>();
  return cudaErrorInvalidValue;
}

int launch_attn_fused(const Dims& d, int B, const int* lens, const int* active, const float* enc_feat,
                      const float* enc_states, const float* dfeat, const float* cov, const float* v, const float* w_c,
                      float* e, float* pctx, float* segstat, int* counters, float* attn, float* ctx, int ctx_ld, int n_sms,
                      cudaStream_t st) {
  AfArgs a;
  a.lens = lens; a.active = active; a.enc_feat = enc_feat; a.enc_states = enc_states; a.dfeat = dfeat;
  a.cov = d.use_cov ? cov : nullptr; a.v = v; a.w_c = w_c; a.e = e; a.pctx = pctx; a.segstat = segstat;
  a.counters = counters; a.attn = attn; a.ctx = ctx; a.ctx_ld = ctx_ld; a.B = B; a.T = d.T; a.nseg_max = attn_fused_nseg(d);
  const int items = B * a.nseg_max;
  const int grid = items < n_sms ? items : n_sms;
  if (d.K == 4) attn_fused_kernel<4><<<grid, AF_NT, AfCfg<4>::SMEM, st>>>(a);
  else attn_fused_kernel<8><<<grid, AF_NT, AfCfg<8>::SMEM, st>>>(a);
  return 1;
}

}  // namespace pgn
