// sm_100a kernels of the pointer-generator decode path: fp32 GEMMs, the persistent cluster BiLSTM, attention, p_gen,
// beam bookkeeping and the all-fp32 vocabulary path (the tcgen05 kernels live in vocab_tc.cu).
// File:line citations are relative to /root/reference/src/main/python/pointer-generator/.
//
// Row convention: decoder row r = b*K + k (article b, beam slot k).  Every per-element reduction
// runs in an order that depends only on that element's own row/article, so an article's result
// does not depend on what else shares the device batch.
#include "kernels.h"

#include <cooperative_groups.h>

#include <cfloat>
#include <cstdlib>
#include <cmath>

namespace pgn {


// ======================================================================================
// Generic fp32 GEMM  C = act([A0|A1] . B^T + bias)
// ======================================================================================
constexpr int GBK = 16;

template <int BM, int BN, int TM, int TN>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
gemm_nt_kernel(GemmArgs g) {
  constexpr int NT = (BM / TM) * (BN / TN);
  __shared__ __align__(16) float As[GBK][BM + 4];
  __shared__ __align__(16) float Bs[GBK][BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid % (BN / TN), ty = tid / (BN / TN);
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int Ktot = g.k0 + g.k1;
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  for (int kt = 0; kt < Ktot; kt += GBK) {
    const bool seg0 = kt < g.k0;
    // A tile: BM rows x 16 k  (BM*4 float4)
    for (int idx = tid; idx < BM * 4; idx += NT) {
      const int row = idx >> 2, k4 = (idx & 3) * 4;
      const int m = m0 + row;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m < g.M) {
        const float* src;
        if (seg0) {
          const long long rr = g.gather0 ? (long long)g.gather0[m] : (long long)m;
          src = g.A0 + rr * g.lda0 + kt + k4;
        } else {
          src = g.A1 + (long long)m * g.lda1 + (kt - g.k0) + k4;
        }
        v = *reinterpret_cast<const float4*>(src);
      }
      As[k4 + 0][row] = v.x; As[k4 + 1][row] = v.y; As[k4 + 2][row] = v.z; As[k4 + 3][row] = v.w;
    }
    for (int idx = tid; idx < BN * 4; idx += NT) {
      const int row = idx >> 2, k4 = (idx & 3) * 4;
      const int n = n0 + row;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (n < g.N) v = *reinterpret_cast<const float4*>(g.B + (long long)n * g.ldb + kt + k4);
      Bs[k4 + 0][row] = v.x; Bs[k4 + 1][row] = v.y; Bs[k4 + 2][row] = v.z; Bs[k4 + 3][row] = v.w;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GBK; ++k) {
      float a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; i += 4) {
        const float4 t = *reinterpret_cast<const float4*>(&As[k][ty * TM + i]);
        a[i] = t.x; a[i + 1] = t.y; a[i + 2] = t.z; a[i + 3] = t.w;
      }
#pragma unroll
      for (int j = 0; j < TN; j += 4) {
        const float4 t = *reinterpret_cast<const float4*>(&Bs[k][tx * TN + j]);
        b[j] = t.x; b[j + 1] = t.y; b[j + 2] = t.z; b[j + 3] = t.w;
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + ty * TM + i;
    if (m >= g.M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int n = n0 + tx * TN + j;
      if (n >= g.N) continue;
      float v = acc[i][j] + (g.bias ? g.bias[n] : 0.f);
      if (g.act == 1) v = fmaxf(v, 0.f);
      g.C[(long long)m * g.ldc + n] = v;
      if (g.C_tf32) {
        uint32_t r;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
        g.C_tf32[(long long)m * g.ldc_tf32 + n] = __uint_as_float(r);
      }
    }
  }
}

// Skinny variant for the per-step decoder GEMMs (M = articles x beam, a few hundred rows): a 32 x 32 output
// tile per CTA, the K loop split over KSPLIT groups of 64 threads that each own every KSPLIT-th k-tile, partial
// sums combined through shared memory in group order (deterministic, independent of the batch).
template <int KSPLIT>
__global__ void __launch_bounds__(64 * KSPLIT)
gemm_skinny_kernel(GemmArgs g) {
  constexpr int BM = 32, BN = 32, LD = 36;
  __shared__ __align__(16) float sm[KSPLIT * 2 * GBK * LD];
  const int tid = threadIdx.x, grp = tid >> 6, t = tid & 63;
  float (*As)[LD] = reinterpret_cast<float (*)[LD]>(sm + grp * 2 * GBK * LD);
  float (*Bs)[LD] = reinterpret_cast<float (*)[LD]>(sm + grp * 2 * GBK * LD + GBK * LD);
  const int tx = t & 7, ty = t >> 3;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int Ktot = g.k0 + g.k1;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  // register-staged software pipeline: the global loads of k-tile i+1 are in flight while tile i is multiplied
  float4 pa[2], pb[2];
  auto fetch = [&](int kt) {
    const bool seg0 = kt < g.k0;
#pragma unroll
    for (int rep = 0; rep < 2; ++rep) {
      const int idx = t + rep * 64, row = idx >> 2, k4 = (idx & 3) * 4;
      const int m = m0 + row, n = n0 + row;
      pa[rep] = make_float4(0.f, 0.f, 0.f, 0.f);
      pb[rep] = pa[rep];
      if (m < g.M) {
        const float* src;
        if (seg0) {
          const long long rr = g.gather0 ? (long long)g.gather0[m] : (long long)m;
          src = g.A0 + rr * g.lda0 + kt + k4;
        } else {
          src = g.A1 + (long long)m * g.lda1 + (kt - g.k0) + k4;
        }
        pa[rep] = *reinterpret_cast<const float4*>(src);
      }
      if (n < g.N) pb[rep] = *reinterpret_cast<const float4*>(g.B + (long long)n * g.ldb + kt + k4);
    }
  };
  int kt = grp * GBK;
  if (kt < Ktot) fetch(kt);
  for (; kt < Ktot; kt += KSPLIT * GBK) {
#pragma unroll
    for (int rep = 0; rep < 2; ++rep) {
      const int idx = t + rep * 64, row = idx >> 2, k4 = (idx & 3) * 4;
      As[k4 + 0][row] = pa[rep].x; As[k4 + 1][row] = pa[rep].y; As[k4 + 2][row] = pa[rep].z; As[k4 + 3][row] = pa[rep].w;
      Bs[k4 + 0][row] = pb[rep].x; Bs[k4 + 1][row] = pb[rep].y; Bs[k4 + 2][row] = pb[rep].z; Bs[k4 + 3][row] = pb[rep].w;
    }
    asm volatile("bar.sync %0, 64;" ::"r"(grp + 1));           // only this group's 64 threads
    if (kt + KSPLIT * GBK < Ktot) fetch(kt + KSPLIT * GBK);
#pragma unroll
    for (int k = 0; k < GBK; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    asm volatile("bar.sync %0, 64;" ::"r"(grp + 1));
  }
  __syncthreads();
  float* red = sm;                                               // [KSPLIT][32][32], aliases the tiles
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) red[grp * 1024 + (ty * 4 + i) * 32 + tx * 4 + j] = acc[i][j];
  __syncthreads();
  for (int o = tid; o < 1024; o += 64 * KSPLIT) {
    const int m = m0 + (o >> 5), n = n0 + (o & 31);
    if (m >= g.M || n >= g.N) continue;
    float v = 0.f;
#pragma unroll
    for (int q = 0; q < KSPLIT; ++q) v += red[q * 1024 + o];
    v += g.bias ? g.bias[n] : 0.f;
    if (g.act == 1) v = fmaxf(v, 0.f);
    g.C[(long long)m * g.ldc + n] = v;
    if (g.C_tf32) {
      uint32_t r;
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
      g.C_tf32[(long long)m * g.ldc_tf32 + n] = __uint_as_float(r);
    }
    if (g.C_pack) {
      const int art = m / g.pack_K, k = m - art * g.pack_K;
      g.C_pack[(((long long)art * (g.pack_K / 2) + (k >> 1)) * g.N + n) * 2 + (k & 1)] = 2.885390081777927f * v;
    }
  }
}
static_assert(8 * 2 * GBK * 36 >= 8 * 1024, "reduction buffer must fit in the tile storage");

int launch_gemm_nt(const GemmArgs& a, cudaStream_t st) {
  if (a.M <= 0 || a.N <= 0) return 0;
  const int ktiles = (a.k0 + a.k1) / GBK;
  if (a.skinny && ktiles >= 8) {
    dim3 grid((a.N + 31) / 32, (a.M + 31) / 32);
    if (a.N <= 512 && ktiles >= 16) gemm_skinny_kernel<8><<<grid, 512, 0, st>>>(a);
    else gemm_skinny_kernel<4><<<grid, 256, 0, st>>>(a);
    return 1;
  }
  const long long work = (long long)a.M * a.N;
  if (work >= 128LL * 128 * 296) {
    dim3 grid((a.N + 127) / 128, (a.M + 127) / 128);
    gemm_nt_kernel<128, 128, 8, 8><<<grid, 256, 0, st>>>(a);
  } else if (work >= 64LL * 64 * 148) {
    dim3 grid((a.N + 63) / 64, (a.M + 63) / 64);
    gemm_nt_kernel<64, 64, 4, 4><<<grid, 256, 0, st>>>(a);
  } else {
    dim3 grid((a.N + 31) / 32, (a.M + 31) / 32);
    gemm_nt_kernel<32, 32, 4, 4><<<grid, 64, 0, st>>>(a);
  }
  return 1;
}

// ======================================================================================
// Encoder LSTM time step (both directions)
// ======================================================================================
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

constexpr int ENC_AG = 8;   // articles per block

__global__ void __launch_bounds__(128)
enc_lstm_step_kernel(int H, int T, int B, int s, const int* __restrict__ lens, const float* __restrict__ zx,
                     const float* __restrict__ kh, const float* __restrict__ h_prev, float* __restrict__ h_next,
                     float* __restrict__ c_state, float* __restrict__ enc_states) {
  extern __shared__ float sm[];
  float* hs = sm;                       // [ENC_AG][H]
  float* zs = sm + ENC_AG * H;          // [ENC_AG][4][32]
  const int tid = threadIdx.x;
  const int chunk = blockIdx.x, dir = blockIdx.y, b0 = blockIdx.z * ENC_AG;
  const int gate = tid >> 5, ul = tid & 31, u = chunk * 32 + ul;
  const int n = gate * H + u;
  const int H4 = 4 * H;
  const float* hp = h_prev + (size_t)dir * B * H;
  for (int i = tid; i < ENC_AG * H; i += 128) {
    const int a = i / H, k = i - a * H;
    hs[i] = (b0 + a < B) ? hp[(size_t)(b0 + a) * H + k] : 0.f;
  }
  float acc[ENC_AG];
#pragma unroll
  for (int a = 0; a < ENC_AG; ++a) {
    const int b = b0 + a;
    int t = -1;
    if (b < B) {
      const int len = lens[b];
      if (s < len) t = dir == 0 ? s : len - 1 - s;      // bw: reverse over the article's own length
    }
    acc[a] = t >= 0 ? zx[((size_t)dir * B * T + (size_t)b * T + t) * H4 + n] : 0.f;
  }
  __syncthreads();
  const float* w = kh + (size_t)dir * H * H4 + n;
  for (int k = 0; k < H; ++k) {
    const float wk = w[(size_t)k * H4];
#pragma unroll
    for (int a = 0; a < ENC_AG; ++a) acc[a] = fmaf(hs[a * H + k], wk, acc[a]);
  }
#pragma unroll
  for (int a = 0; a < ENC_AG; ++a) zs[(a * 4 + gate) * 32 + ul] = acc[a];
  __syncthreads();
  for (int i = tid; i < ENC_AG * 32; i += 128) {
    const int a = i >> 5, l = i & 31, b = b0 + a, uu = chunk * 32 + l;
    if (b >= B) continue;
    const size_t si = (size_t)dir * B * H + (size_t)b * H + uu;
    // recompute t for article a (tpos[] is indexed by a compile-time-unrolled a only)
    const int len = lens[b];
    if (s < len) {
      const int t = dir == 0 ? s : len - 1 - s;
      const float zi = zs[(a * 4 + 0) * 32 + l], zj = zs[(a * 4 + 1) * 32 + l];
      const float zf = zs[(a * 4 + 2) * 32 + l], zo = zs[(a * 4 + 3) * 32 + l];
      const float c = sigmoidf_(zf + 1.0f) * c_state[si] + sigmoidf_(zi) * tanhf(zj);   // LSTMCell, forget_bias 1
      const float h = sigmoidf_(zo) * tanhf(c);
      c_state[si] = c;
      h_next[si] = h;
      enc_states[((size_t)b * T + t) * (2 * H) + dir * H + uu] = h;                       // model.py:93
    } else {
      h_next[si] = hs[a * H + uu];                      // past the end: state frozen (dynamic_rnn)
    }
  }
}

int launch_enc_lstm_step(const Dims& d, int B, int s, const int* lens, const float* zx, const float* kh,
                         const float* h_prev, float* h_next, float* c_state, float* enc_states,
                         cudaStream_t st) {
  dim3 grid(d.H / 32, 2, (B + ENC_AG - 1) / ENC_AG);
  const size_t smem = (size_t)(ENC_AG * d.H + ENC_AG * 4 * 32) * sizeof(float);
  enc_lstm_step_kernel<<<grid, 128, smem, st>>>(d.H, d.T, B, s, lens, zx, kh, h_prev, h_next, c_state, enc_states);
  return 1;
}

// --------------------------------------------------------------------------------------
// Persistent encoder recurrence for hidden_dim = 256 (model.py:90-93): one launch for all time steps.
// A cluster of 8 CTAs owns one (direction, group of 8 articles); CTA c of the cluster holds the recurrent
// weights of hidden units [32c, 32c+32) (4 gates x 32 = 128 columns, 128 KB fp32) in shared memory for the
// whole sequence, computes their gates for the 8 articles, and pushes the 32 new h values into every CTA's
// shared memory through DSMEM; one cluster barrier per time step.  c stays in registers.
// --------------------------------------------------------------------------------------
namespace cgx = cooperative_groups;
// packed fp32 pairs (FFMA2 / FMUL2 / FADD2): each half is an ordinary IEEE fp32 operation
__device__ __forceinline__ unsigned long long pk2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk2(unsigned long long v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ unsigned long long fma2_(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ unsigned long long mul2_(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ unsigned long long add2_(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

constexpr int PE_H = 256, PE_THREADS = 512;
template <int AG> struct PeCfg {                      // AG articles per cluster
  static constexpr int AGP = (AG + 3) & ~3;           // row pitch of the h buffers (float4 loads)
  static constexpr size_t SMEM = (size_t)(PE_H * 128 + 2 * PE_H * AGP + 8 * AG * 128 + 2 * 32 * AGP) * sizeof(float) + 64;
  static constexpr uint32_t H_BYTES = PE_H * AGP * sizeof(float);  // one full h(s) of the group (padded rows)
};

__device__ __forceinline__ uint32_t pe_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t pe_mapa(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// DSMEM store that signals the destination CTA's mbarrier with the bytes written (no cluster-wide fence)
__device__ __forceinline__ void pe_st_async(uint32_t raddr, float v, uint32_t rbar) {
  asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];"
               ::"r"(raddr), "r"(__float_as_uint(v)), "r"(rbar) : "memory");
}
__device__ __forceinline__ void pe_bar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok = 0, spins = 0;
  const uint32_t a = pe_smem_u32(bar);
  while (!ok) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    if (!ok && ++spins > (1u << 26)) __trap();
  }
}

// Persistent BiLSTM recurrence (model.py:76-93): one cluster of 8 CTAs per (direction, group of AG articles) runs all
// time steps.  CTA c keeps the recurrent weights of hidden units [32c, 32c+32) (4 gates = 128 columns, 128 KB) in
// shared memory for the whole kernel, computes its slice of h.Kh with the K range split over 8 thread groups, adds
// the partial sums in fixed order, applies the gates and sends its 32 x AG new h values to every CTA of the cluster
// with st.async, which also signals the receiver's mbarrier: no cluster-wide barrier per step.
// AG is chosen so that ALL clusters are co-resident: a B200 runs at most 15 clusters of 8 CTAs with this footprint,
// and a 16th would cost a whole second wave (tools/micro/cluster_occ.cu).  Every article's arithmetic is the same
// for every AG.
#ifdef PGEN_LIVE_PE
__device__ unsigned int live_pe[16];                 // entry / exit counters (pgen_debug_where)
#define LIVE_PE(i) atomicAdd(&live_pe[i], 1u)
#else
#define LIVE_PE(i) ((void)0)
#endif

template <int AG>
__global__ void __cluster_dims__(8, 1, 1) __launch_bounds__(PE_THREADS, 1)
enc_lstm_persistent_kernel(int T, int B, const int* __restrict__ lens, const float* __restrict__ zx,
                           const int* __restrict__ ids, int V,
                           const float* __restrict__ kh, float* __restrict__ h_fin, float* __restrict__ c_fin,
                           float* __restrict__ enc_states) {
  constexpr int AGP = PeCfg<AG>::AGP;
  constexpr uint32_t H_BYTES = PeCfg<AG>::H_BYTES;
  extern __shared__ __align__(16) float sm[];
  float* Ws = sm;                                  // [256 k][128 local cols]
  float* hs = Ws + PE_H * 128;                     // [2][256 k][AGP articles]
  float* red = hs + 2 * PE_H * AGP;                // [8 kq][AG articles][128 cols]
  float* hstage = red + 8 * AG * 128;              // [2][32 units][AGP] this CTA's new h, source of the bulk copies
  uint64_t* hbar = reinterpret_cast<uint64_t*>(hstage + 2 * 32 * AGP);   // [2] "h buffer b is complete"
  cgx::cluster_group cluster = cgx::this_cluster();
  const int c = (int)cluster.block_rank();         // which 32 hidden units
  const int dir = blockIdx.y, b0 = blockIdx.z * AG;
  const int tid = threadIdx.x;
  constexpr int H4 = 4 * PE_H;
  if (tid == 0) LIVE_PE(0);

  // recurrent weights of this CTA's 128 gate columns, and h(-1) = 0
  const float* khd = kh + (size_t)dir * PE_H * H4;
  for (int i = tid; i < PE_H * 128; i += PE_THREADS) {
    const int k = i >> 7, lc = i & 127, g = lc >> 5, ul = lc & 31;
    Ws[i] = khd[(size_t)k * H4 + g * PE_H + 32 * c + ul];
  }
  for (int i = tid; i < 2 * PE_H * AGP; i += PE_THREADS) hs[i] = 0.f;
  for (int i = tid; i < 2 * 32 * AGP; i += PE_THREADS) hstage[i] = 0.f;
  if (tid == 0) {
    for (int b = 0; b < 2; ++b) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(pe_smem_u32(&hbar[b])));
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pe_smem_u32(&hbar[b])), "r"(H_BYTES) : "memory");
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  int grp_max = 0;
  for (int a = 0; a < AG; ++a) if (b0 + a < B) grp_max = max(grp_max, lens[b0 + a]);
  // roles
  const int kq = tid >> 6, cp = tid & 63;          // mat-vec: k range [32kq, 32kq+32), columns cp and cp+64
  const int ful = tid & 31, fa = tid >> 5;         // gate math (threads < 32 AG): unit ful, article fa
  const bool gate_thread = fa < AG;
  const int fb = b0 + fa;
  const int flen = (gate_thread && fb < B) ? lens[fb] : 0;
  float c_reg = 0.f, h_reg = 0.f;
  int id_cur = (ids && flen > 0) ? __ldg(ids + (size_t)fb * T + (dir == 0 ? 0 : flen - 1)) : 0;
  // thread rk < 8 sends this CTA's 32 x AGP block of new h to CTA rk of the cluster with ONE bulk copy per step
  // (cp.async.bulk shared::cta -> shared::cluster), which also signals that CTA's mbarrier with the bytes written
  const uint32_t r_dst = pe_mapa(pe_smem_u32(hs + 32 * c * AGP), tid & 7);
  const uint32_t r_bar = pe_mapa(pe_smem_u32(&hbar[0]), tid & 7);
  cluster.sync();

  for (int s = 0; s < grp_max; ++s) {
    const int cur = s & 1;
    // prefetch this step's input projection (x.Kx + b) for the gate-math threads
    float zi = 0.f, zj = 0.f, zf = 0.f, zo = 0.f;
    int t = -1;
    if (gate_thread && s < flen) {
      t = dir == 0 ? s : flen - 1 - s;             // backward: reversed over the article's own length
      // ids != 0: zx is the per-word table [2][V][4H] (x.Kx + b of every vocabulary word, computed once per model)
      // and the row is the token's; otherwise zx is this batch's projection [2][B][T][4H]
      const float* zp = ids ? zx + ((size_t)dir * V + id_cur) * H4 + 32 * c + ful
                            : zx + ((size_t)dir * B * T + (size_t)fb * T + t) * H4 + 32 * c + ful;
      zi = __ldg(zp); zj = __ldg(zp + PE_H); zf = __ldg(zp + 2 * PE_H); zo = __ldg(zp + 3 * PE_H);
      if (ids && s + 1 < flen) id_cur = __ldg(ids + (size_t)fb * T + (dir == 0 ? s + 1 : flen - 2 - s));   // next step's token
    }
    if (s > 0) {                                   // h(s-1) from all 8 CTAs has landed in hs[cur]
      pe_bar_wait(&hbar[cur], ((s - 1) >> 1) & 1);
      if (tid == 0 && s + 1 < grp_max)             // arm the same barrier for h(s+1), which reuses this buffer
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pe_smem_u32(&hbar[cur])), "r"(H_BYTES) : "memory");
    }
    // partial h . Kh over this thread's k range
    // packed fp32 pairs over two articles: acc0p[q] = (article 2q, article 2q+1) of column cp, acc1p of column cp + 64
    unsigned long long acc0p[AGP / 2], acc1p[AGP / 2];
#pragma unroll
    for (int a = 0; a < AGP / 2; ++a) { acc0p[a] = 0ull; acc1p[a] = 0ull; }
    const float* hcur = hs + cur * PE_H * AGP;
#pragma unroll 4
    for (int kk = 0; kk < 32; ++kk) {
      const int k = kq * 32 + kk;
      const float w0 = Ws[k * 128 + cp], w1 = Ws[k * 128 + cp + 64];
      const unsigned long long w0p = pk2(w0, w0), w1p = pk2(w1, w1);
#pragma unroll
      for (int q = 0; q < AGP / 4; ++q) {
        const ulonglong2 h4 = *reinterpret_cast<const ulonglong2*>(hcur + k * AGP + 4 * q);
        acc0p[2 * q] = fma2_(h4.x, w0p, acc0p[2 * q]); acc1p[2 * q] = fma2_(h4.x, w1p, acc1p[2 * q]);
        if (4 * q + 2 < AG) { acc0p[2 * q + 1] = fma2_(h4.y, w0p, acc0p[2 * q + 1]); acc1p[2 * q + 1] = fma2_(h4.y, w1p, acc1p[2 * q + 1]); }
      }
    }
    float acc0[AGP], acc1[AGP];
#pragma unroll
    for (int a = 0; a < AGP / 2; ++a) { upk2(acc0p[a], acc0[2 * a], acc0[2 * a + 1]); upk2(acc1p[a], acc1[2 * a], acc1[2 * a + 1]); }
    // The partial sums of step s overwrite those of step s-1: every thread that read them has since issued its
    // st.async, and nobody gets here before ALL of those have landed (the wait above).
#pragma unroll
    for (int a = 0; a < AG; ++a) {
      red[(kq * AG + a) * 128 + cp] = acc0[a];
      red[(kq * AG + a) * 128 + cp + 64] = acc1[a];
    }
    __syncthreads();
    if (gate_thread) {
      float hn = h_reg;
      if (t >= 0) {
        float gi = zi, gj = zj, gf = zf, go = zo;
#pragma unroll
        for (int q = 0; q < 8; ++q) {              // fixed order: deterministic
          const float* r = red + (q * AG + fa) * 128 + ful;
          gi += r[0]; gj += r[32]; gf += r[64]; go += r[96];
        }
        c_reg = sigmoidf_(gf + 1.0f) * c_reg + sigmoidf_(gi) * tanhf(gj);   // LSTMCell, forget_bias 1
        hn = sigmoidf_(go) * tanhf(c_reg);
        h_reg = hn;
        enc_states[((size_t)fb * T + t) * (2 * PE_H) + dir * PE_H + 32 * c + ful] = hn;   // model.py:93
      }
      // new h (or the frozen one past the article's end) into this step's staging buffer
      if (s + 1 < grp_max) {
        hstage[(cur ^ 1) * 32 * AGP + ful * AGP + fa] = hn;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> async-proxy reads
      }
    }
    if (s + 1 < grp_max && tid < 32 * AG) {
      asm volatile("bar.sync 1, %0;" ::"n"(32 * AG) : "memory");          // the gate warps only
      if (tid < 8) {
        // The staging buffer written two steps ago is free: every CTA has since passed a wait that needed that copy.
        const uint32_t src = pe_smem_u32(hstage + (cur ^ 1) * 32 * AGP);
        const uint32_t boff = (uint32_t)(cur ^ 1) * (uint32_t)(PE_H * AGP * sizeof(float)), bar_off = (uint32_t)(cur ^ 1) * 8u;
        asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(r_dst + boff), "r"(src), "n"(32 * AGP * 4), "r"(r_bar + bar_off) : "memory");
      }
    }
  }
  if (tid == 0) LIVE_PE(2);
  cluster.sync();                                  // nobody leaves while remote stores may still target it
  if (tid == 0) LIVE_PE(1);
  if (gate_thread && fb < B) {
    h_fin[(size_t)dir * B * PE_H + (size_t)fb * PE_H + 32 * c + ful] = h_reg;
    c_fin[(size_t)dir * B * PE_H + (size_t)fb * PE_H + 32 * c + ful] = c_reg;
  }
}

template <int AG>
static int pe_launch(const Dims& d, int B, const int* lens, const float* zx, const int* ids, const float* kh, float* h_fin,
                     float* c_fin, float* enc_states, cudaStream_t st) {
  static PerDeviceOnce once;
  once.ensure(1, [] { cudaFuncSetAttribute(enc_lstm_persistent_kernel<AG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PeCfg<AG>::SMEM); });
  dim3 grid(8, 2, (B + AG - 1) / AG);
  enc_lstm_persistent_kernel<AG><<<grid, PE_THREADS, PeCfg<AG>::SMEM, st>>>(d.T, B, lens, zx, ids, d.V, kh, h_fin, c_fin, enc_states);
  return 1;
}

// How many clusters of 8 CTAs with this kernel's footprint run at once on the current device.
static int pe_max_clusters() {
  static int cached[64] = {0};
  static std::mutex mu;
  std::lock_guard<std::mutex> lk(mu);
  int dev = 0;
  cudaGetDevice(&dev);
  if (cached[dev & 63] > 0) return cached[dev & 63];
  cudaFuncSetAttribute(enc_lstm_persistent_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PeCfg<8>::SMEM);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(8, 2, 8); cfg.blockDim = dim3(PE_THREADS); cfg.dynamicSmemBytes = PeCfg<8>::SMEM;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 8; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, enc_lstm_persistent_kernel<8>, &cfg) != cudaSuccess || n < 2) { cudaGetLastError(); n = 14; }
  cached[dev & 63] = n;
  return n;
}

int launch_enc_lstm_persistent(const Dims& d, int B, const int* lens, const float* zx, const int* ids, const float* kh,
                               float* h_fin, float* c_fin, float* enc_states, cudaStream_t st) {
  // the smallest group size whose 2 x ceil(B / AG) clusters are all co-resident
  const int per_dir = pe_max_clusters() / 2;
  static const char* force = getenv("PGEN_LSTM_AG");
  const int forced = force ? atoi(force) : 0;
  if (forced == 8 || (!forced && (B + 7) / 8 <= per_dir)) return pe_launch<8>(d, B, lens, zx, ids, kh, h_fin, c_fin, enc_states, st);
  if (forced == 10 || (!forced && (B + 9) / 10 <= per_dir)) return pe_launch<10>(d, B, lens, zx, ids, kh, h_fin, c_fin, enc_states, st);
  return pe_launch<12>(d, B, lens, zx, ids, kh, h_fin, c_fin, enc_states, st);     // larger batches: more than one wave
}

// ======================================================================================
// Decoder step: gather by parent, LSTM pointwise
// ======================================================================================
__global__ void __launch_bounds__(128)
gather_step_kernel(Dims d, StepBufs b) {
  const int r = blockIdx.x;
  const int art = r / d.K;
  if (!b.active[art]) return;
  const int p = b.sel_parent[r];
  int tok = b.sel_token[r];
  if (tok >= d.V) tok = 0;                                           // beam_search.py:112-113  OOV -> [UNK]
  const int rp = art * d.K + p;
  const float* so = b.state_old + (size_t)rp * d.SW;
  float* xin = b.xin + (size_t)r * (d.E + d.A);
  const float* er = b.emb + (size_t)tok * d.E;
  const int len = b.lens[art];
  if (((d.E | d.A | d.H | d.T | d.SW) & 3) == 0) {
    // Everything is a multiple of 4 floats: the row is a list of 128-bit copies, and a thread requests ALL of its
    // chunks before it stores the first (one memory round trip instead of one per loop iteration).
    const int nE = d.E / 4, nA = d.A / 4, nH = d.H / 4, nT = d.use_cov ? (len + 3) / 4 : 0;
    const int total = nE + nA + 2 * nH + nT;
    const float4* co = reinterpret_cast<const float4*>(b.cov_old + (size_t)rp * d.T);
    const float4* ao = reinterpret_cast<const float4*>(b.attn_old + (size_t)rp * d.T);
    constexpr int MAXC = 6;                                          // 128 threads x 6 chunks x 4 floats per pass
    for (int base = 0; base < total; base += 128 * MAXC) {
      float4 v[MAXC], w[MAXC];
#pragma unroll
      for (int q = 0; q < MAXC; ++q) {
        const int i = base + q * 128 + threadIdx.x;
        w[q] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < nE) v[q] = __ldg(reinterpret_cast<const float4*>(er) + i);
        else if (i < nE + nA + 2 * nH) v[q] = *(reinterpret_cast<const float4*>(so) + (i - nE));   // ctx | c | h are contiguous
        else if (i < total) { v[q] = co[i - nE - nA - 2 * nH]; w[q] = ao[i - nE - nA - 2 * nH]; }
      }
#pragma unroll
      for (int q = 0; q < MAXC; ++q) {
        const int i = base + q * 128 + threadIdx.x;
        if (i < nE + nA) reinterpret_cast<float4*>(xin)[i] = v[q];                                  // model.py:214 | previous context
        else if (i < nE + nA + nH) reinterpret_cast<float4*>(b.cprev + (size_t)r * d.H)[i - nE - nA] = v[q];
        else if (i < nE + nA + 2 * nH) reinterpret_cast<float4*>(b.hprev + (size_t)r * d.H)[i - nE - nA - nH] = v[q];
        else if (i < total) {                                                                      // attention_decoder.py:114
          const int t4 = i - nE - nA - 2 * nH;
          reinterpret_cast<float4*>(b.cov_new + (size_t)r * d.T)[t4] =
              make_float4(v[q].x + w[q].x, v[q].y + w[q].y, v[q].z + w[q].z, v[q].w + w[q].w);
        }
      }
    }
    return;
  }
  for (int i = threadIdx.x; i < d.E; i += 128) xin[i] = er[i];       // model.py:214
  for (int i = threadIdx.x; i < d.A; i += 128) xin[d.E + i] = so[i]; // previous context
  for (int i = threadIdx.x; i < d.H; i += 128) {
    b.cprev[(size_t)r * d.H + i] = so[d.A + i];
    b.hprev[(size_t)r * d.H + i] = so[d.A + d.H + i];
  }
  if (d.use_cov) {                                                   // attention_decoder.py:114
    const float* co = b.cov_old + (size_t)rp * d.T;
    const float* ao = b.attn_old + (size_t)rp * d.T;
    float* cn = b.cov_new + (size_t)r * d.T;
    for (int t = threadIdx.x; t < len; t += 128) cn[t] = co[t] + ao[t];
  }
}

int launch_gather_step(const Dims& d, int B, const StepBufs& b, cudaStream_t st) {
  gather_step_kernel<<<B * d.K, 128, 0, st>>>(d, b);
  return 1;
}

__global__ void lstm_pointwise_kernel(int H, int A, int SW, int K, int R, const float* __restrict__ z,
                                      const float* __restrict__ cprev, float* __restrict__ state_new,
                                      const int* __restrict__ active) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R * H) return;
  const int r = i / H, u = i - r * H;
  if (!active[r / K]) return;
  const float* zr = z + (size_t)r * 4 * H;
  const float zi = zr[u], zj = zr[H + u], zf = zr[2 * H + u], zo = zr[3 * H + u];
  const float c = sigmoidf_(zf + 1.0f) * cprev[i] + sigmoidf_(zi) * tanhf(zj);
  const float h = sigmoidf_(zo) * tanhf(c);
  state_new[(size_t)r * SW + A + u] = c;
  state_new[(size_t)r * SW + A + H + u] = h;
}

int launch_lstm_pointwise(const Dims& d, int R, const float* z, const float* cprev, float* state_new,
                          const int* active, cudaStream_t st) {
  const int n = R * d.H;
  lstm_pointwise_kernel<<<(n + 255) / 256, 256, 0, st>>>(d.H, d.A, d.SW, d.K, R, z, cprev, state_new, active);
  return 1;
}

// ======================================================================================
// Attention: energies, then softmax + context
// ======================================================================================
constexpr int ATT_TCH = 16;     // encoder positions per block (2 per warp)

// tanh(x) = 1 - 2 / (1 + e^{2x}) on the special-function unit (ex2.approx + rcp.approx): absolute error ~2e-7,
// saturates to +-1 for large |x|.  tanhf() costs ~4x the instructions and made the energy kernel issue-bound.
__device__ __forceinline__ float tanh_mufu(float x) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * 2.885390081777927f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return fmaf(-2.0f, r, 1.0f);
}

__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <int KMAX, bool FULL>                     // FULL: all KMAX rows of the article are computed (rpa == KMAX)
__global__ void __launch_bounds__(256)
attn_energy_kernel(Dims d, int rpa_arg, const int* __restrict__ lens, const int* __restrict__ active,
                   const float* __restrict__ enc_feat, const float* __restrict__ dfeat,
                   const float* __restrict__ cov, const float* __restrict__ v, const float* __restrict__ w_c,
                   float* __restrict__ e) {
  extern __shared__ __align__(16) float sm[];
  const int art = blockIdx.y;
  if (!active[art]) return;
  const int len = lens[art];
  const int t0 = blockIdx.x * ATT_TCH;
  if (t0 >= len) return;
  const int A = d.A;
  const int rpa = FULL ? KMAX : rpa_arg;
  constexpr float kS = 2.885390081777927f;          // 2 log2(e): everything below is pre-scaled, e^{2x} = 2^{kS x}
  float* ds = sm;                 // [rpa][A]   kS * decoder features
  float* vs = ds + KMAX * A;      // [A]
  float* ws = vs + A;             // [A]
  for (int k = 0; k < rpa; ++k) {                       // no i / A here: a runtime integer division per element
    const float* src = dfeat + (size_t)(art * d.K + k) * A;           // costs more than the loads it indexes
    for (int a = threadIdx.x; a < A; a += 256) ds[k * A + a] = kS * src[a];
  }
  for (int i = threadIdx.x; i < A; i += 256) { vs[i] = -2.0f * v[i]; ws[i] = d.use_cov ? kS * w_c[i] : 0.f; }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // sum of v over a: e = sum_a v_a tanh(x_a) = sum_a v_a - 2 sum_a v_a / (1 + exp(2 x_a))
  float vsum = 0.f;
  for (int a = lane; a < A; a += 32) vsum += vs[a];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) vsum += __shfl_xor_sync(0xffffffffu, vsum, o);
  vsum *= -0.5f;
  for (int tt = warp; tt < ATT_TCH; tt += 8) {
    const int t = t0 + tt;
    if (t >= len) break;
    float cv[KMAX], acc[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      acc[k] = 0.f;
      cv[k] = (d.use_cov && k < rpa) ? cov[(size_t)(art * d.K + k) * d.T + t] : 0.f;
    }
    const float4* f4 = reinterpret_cast<const float4*>(enc_feat + ((size_t)art * d.T + t) * A);
#pragma unroll 4
    for (int a4 = lane; a4 < A / 4; a4 += 32) {
      const float4 f = __ldg(f4 + a4);
      const float4 vv = *reinterpret_cast<const float4*>(vs + a4 * 4);      // -2 v
      const float4 ww = *reinterpret_cast<const float4*>(ws + a4 * 4);      // kS w_c
#pragma unroll
      for (int k = 0; k < KMAX; ++k) {
        if (FULL || k < rpa) {
          const float4 dd = *reinterpret_cast<const float4*>(ds + k * A + a4 * 4);
          // tanh(x) = 1 - 2 / (1 + e^{2x}).  The four reciprocals share ONE rcp: r = 1 / (A0 A1 A2 A3), 1/A0 = r A1 A2 A3, ...
          // x is clamped at 10 (tanh(10) rounds to 1.0f), so A_i <= 1 + e^20 and the product stays below 6e34.
          const float y0 = fminf(fmaf(cv[k], ww.x, fmaf(f.x, kS, dd.x)), 10.f * kS);
          const float y1 = fminf(fmaf(cv[k], ww.y, fmaf(f.y, kS, dd.y)), 10.f * kS);
          const float y2 = fminf(fmaf(cv[k], ww.z, fmaf(f.z, kS, dd.z)), 10.f * kS);
          const float y3 = fminf(fmaf(cv[k], ww.w, fmaf(f.w, kS, dd.w)), 10.f * kS);
          const float A0 = 1.0f + ex2_approx(y0), A1 = 1.0f + ex2_approx(y1);
          const float A2 = 1.0f + ex2_approx(y2), A3 = 1.0f + ex2_approx(y3);
          const float p01 = A0 * A1, p23 = A2 * A3;
          const float r = rcp_approx(p01 * p23);
          const float r01 = r * p23, r23 = r * p01;                          // 1 / (A0 A1), 1 / (A2 A3)
          float s = acc[k];
          s = fmaf(vv.x, r01 * A1, s); s = fmaf(vv.y, r01 * A0, s);
          s = fmaf(vv.z, r23 * A3, s); s = fmaf(vv.w, r23 * A2, s);
          acc[k] = s;
        }
      }
    }
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      if (k < rpa) {
        float s = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) e[(size_t)(art * d.K + k) * d.T + t] = vsum + s;
      }
    }
  }
}

// All-rows variant (rpa == KMAX, KMAX even) on packed fp32 pairs: the two halves of every FFMA2 / FMUL2 / FADD2 are
// the same encoder feature for two beam rows of the article, so the arithmetic of a row is exactly that of the scalar
// kernel above at half the issue slots.  Lane l owns elements l, l + 32, l + 64, l + 96 of every 128-element chunk
// (they share the reciprocal); the decoder features sit in shared memory as [a][row], one 128-bit read per element.
template <int KMAX, int ACH>                       // A = 128 ACH attention features
__global__ void __launch_bounds__(256, 5)
attn_energy_pair_kernel(Dims d, const int* __restrict__ lens, const int* __restrict__ active,
                        const float* __restrict__ enc_feat, const float* __restrict__ dfeat,
                        const float* __restrict__ cov, const float* __restrict__ v, const float* __restrict__ w_c,
                        float* __restrict__ e) {
  static_assert(KMAX % 2 == 0, "rows are processed in pairs");
  constexpr int NP = KMAX / 2;
  extern __shared__ __align__(16) float sm[];
  const int art = blockIdx.y;
  if (!active[art]) return;
  const int len = lens[art];
  const int t0 = blockIdx.x * ATT_TCH;
  if (t0 >= len) return;
  constexpr int A = 128 * ACH;
  constexpr float kS = 2.885390081777927f;          // 2 log2(e): operands are pre-scaled, e^{2x} = 2^{kS x}
  float* ds = sm;                 // [A][KMAX]  kS * decoder features, row fastest
  float* vs = ds + KMAX * A;      // [A]        -2 v
  float* ws = vs + A;             // [A]        kS w_c
  for (int k = 0; k < KMAX; ++k) {
    const float* src = dfeat + (size_t)(art * d.K + k) * A;
    for (int a = threadIdx.x; a < A; a += 256) ds[a * KMAX + k] = kS * src[a];
  }
  for (int i = threadIdx.x; i < A; i += 256) { vs[i] = -2.0f * v[i]; ws[i] = d.use_cov ? kS * w_c[i] : 0.f; }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float vsum = 0.f;
  for (int a = lane; a < A; a += 32) vsum += vs[a];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) vsum += __shfl_xor_sync(0xffffffffu, vsum, o);
  vsum *= -0.5f;                                    // sum of v: e = sum_a v_a - 2 sum_a v_a / (1 + e^{2 x_a})
  const unsigned long long kS2 = pk2(kS, kS), one2 = pk2(1.0f, 1.0f);
  for (int tt = warp; tt < ATT_TCH; tt += 8) {
    const int t = t0 + tt;
    if (t >= len) break;
    unsigned long long cv2[NP], acc2[NP];
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      acc2[p] = 0ull;
      cv2[p] = d.use_cov ? pk2(cov[(size_t)(art * d.K + 2 * p) * d.T + t], cov[(size_t)(art * d.K + 2 * p + 1) * d.T + t]) : 0ull;
    }
    const float* fr = enc_feat + ((size_t)art * d.T + t) * A + lane;
    float f[4 * ACH];                               // the whole feature row of this position is requested up front
#pragma unroll
    for (int i = 0; i < 4 * ACH; ++i) f[i] = __ldg(fr + 32 * i);
#pragma unroll
    for (int c = 0; c < ACH; ++c) {
      unsigned long long vv2[4], ww2[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float vv = vs[c * 128 + lane + 32 * i], ww = ws[c * 128 + lane + 32 * i];
        vv2[i] = pk2(vv, vv); ww2[i] = pk2(ww, ww);
      }
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        unsigned long long Aq[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const unsigned long long dd = *reinterpret_cast<const unsigned long long*>(ds + (c * 128 + lane + 32 * i) * KMAX + 2 * p);
          const unsigned long long y = fma2_(cv2[p], ww2[i], fma2_(pk2(f[4 * c + i], f[4 * c + i]), kS2, dd));
          float y0, y1;
          upk2(y, y0, y1);
          // x is clamped at 10 (tanh(10) rounds to 1.0f): 1 + e^{2x} <= 1 + e^20 and the product of four stays below 6e34
          Aq[i] = add2_(one2, pk2(ex2_approx(fminf(y0, 10.f * kS)), ex2_approx(fminf(y1, 10.f * kS))));
        }
        // the four reciprocals of a row share ONE rcp: r = 1 / (A0 A1 A2 A3), 1/A0 = r A1 A2 A3, ...
        const unsigned long long p01 = mul2_(Aq[0], Aq[1]), p23 = mul2_(Aq[2], Aq[3]);
        float q0, q1;
        upk2(mul2_(p01, p23), q0, q1);
        const unsigned long long r = pk2(rcp_approx(q0), rcp_approx(q1));
        const unsigned long long r01 = mul2_(r, p23), r23 = mul2_(r, p01);
        unsigned long long s2 = acc2[p];
        s2 = fma2_(vv2[0], mul2_(r01, Aq[1]), s2);
        s2 = fma2_(vv2[1], mul2_(r01, Aq[0]), s2);
        s2 = fma2_(vv2[2], mul2_(r23, Aq[3]), s2);
        s2 = fma2_(vv2[3], mul2_(r23, Aq[2]), s2);
        acc2[p] = s2;
      }
    }
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      float s0, s1;
      upk2(acc2[p], s0, s1);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o); }
      if (lane == 0) {
        e[(size_t)(art * d.K + 2 * p) * d.T + t] = vsum + s0;
        e[(size_t)(art * d.K + 2 * p + 1) * d.T + t] = vsum + s1;
      }
    }
  }
}

int launch_attn_energy(const Dims& d, int B, int rpa, const int* lens, const int* active,
                       const float* enc_feat, const float* dfeat, const float* cov, const float* v,
                       const float* w_c, float* e, cudaStream_t st) {
  dim3 grid((d.T + ATT_TCH - 1) / ATT_TCH, B);
  if (d.K <= 4) {
    const size_t smem = (size_t)(4 * d.A + 2 * d.A) * sizeof(float);
    if (rpa == 4 && d.K == 4 && d.A == 512) attn_energy_pair_kernel<4, 4><<<grid, 256, smem, st>>>(d, lens, active, enc_feat, dfeat, cov, v, w_c, e);
    else attn_energy_kernel<4, false><<<grid, 256, smem, st>>>(d, rpa, lens, active, enc_feat, dfeat, cov, v, w_c, e);
  } else if (d.K <= 8) {
    const size_t smem = (size_t)(8 * d.A + 2 * d.A) * sizeof(float);
    if (rpa == 8 && d.K == 8 && d.A == 512) attn_energy_pair_kernel<8, 4><<<grid, 256, smem, st>>>(d, lens, active, enc_feat, dfeat, cov, v, w_c, e);
    else attn_energy_kernel<8, false><<<grid, 256, smem, st>>>(d, rpa, lens, active, enc_feat, dfeat, cov, v, w_c, e);
  } else {
    const size_t smem = (size_t)(16 * d.A + 2 * d.A) * sizeof(float);
    static PerDeviceOnce once;
    once.ensure(1, [] { cudaFuncSetAttribute(attn_energy_kernel<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024); });
    attn_energy_kernel<16, false><<<grid, 256, smem, st>>>(d, rpa, lens, active, enc_feat, dfeat, cov, v, w_c, e);
  }
  return 1;
}

constexpr int CTX_COLS = 64;    // context columns per block

template <int KMAX>
__global__ void __launch_bounds__(256)
attn_context_kernel(Dims d, int rpa, const int* __restrict__ lens, const int* __restrict__ active,
                    const float* __restrict__ enc_states, const float* __restrict__ e,
                    float* __restrict__ attn, float* __restrict__ state) {
  extern __shared__ __align__(16) float sm[];
  const int art = blockIdx.y;
  if (!active[art]) return;
  const int len = lens[art];
  const int T = d.T, A = d.A;
  float* as = sm;                         // [KMAX][T]
  float* red = sm + KMAX * T;             // [8][KMAX][CTX_COLS]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // The encoder-state loads do not depend on the softmax: put the first four per thread in flight now, so that the
  // HBM latency is paid while the softmax below is computed.
  const int c4 = threadIdx.x & 15, tl = threadIdx.x >> 4;             // 16 float4 columns x 16 t-lanes
  const int col = blockIdx.x * CTX_COLS + c4 * 4;
  const float* hb = enc_states + (size_t)art * T * A + col;
  float4 pre[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int t = tl + 16 * u;
    pre[u] = t < len ? __ldg(reinterpret_cast<const float4*>(hb + (size_t)t * A)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // masked softmax per row (attention_decoder.py:96-101); mask == (t < len)
  for (int k = warp; k < rpa; k += 8) {
    const float* er = e + (size_t)(art * d.K + k) * T;
    float m = -FLT_MAX;
    for (int t = lane; t < len; t += 32) m = fmaxf(m, er[t]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int t = lane; t < len; t += 32) { const float x = expf(er[t] - m); as[k * T + t] = x; s += x; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    float s2 = 0.f;
    for (int t = lane; t < len; t += 32) { const float x = as[k * T + t] / s; as[k * T + t] = x; s2 += x; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    for (int t = lane; t < len; t += 32) {
      const float x = as[k * T + t] / s2;                             // re-normalise (:100-101)
      as[k * T + t] = x;
      if (blockIdx.x == 0) attn[(size_t)(art * d.K + k) * T + t] = x;
    }
    if (blockIdx.x == 0)                                              // the masked distribution is exactly 0 past the article (:99)
      for (int t = len + lane; t < T; t += 32) attn[(size_t)(art * d.K + k) * T + t] = 0.f;
  }
  __syncthreads();
  // context = sum_t attn[t] * h_t   (:126-127)
  float4 acc[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; ++k) acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  // rounds of 4 positions per thread (t = tl + 16 u + 64 round); the loads of the next round are in flight while this
  // round is accumulated, so the memory latency is paid once per thread instead of once per round
  float4 cur[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) cur[u] = pre[u];
  for (int t = tl; t < len; t += 64) {
    float4 nxt[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int tn = t + 64 + 16 * u;
      nxt[u] = tn < len ? __ldg(reinterpret_cast<const float4*>(hb + (size_t)tn * A)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int tu = t + 16 * u;
      if (tu < len) {
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
          if (k < rpa) {
            const float a = as[k * T + tu];
            acc[k].x = fmaf(a, cur[u].x, acc[k].x); acc[k].y = fmaf(a, cur[u].y, acc[k].y);
            acc[k].z = fmaf(a, cur[u].z, acc[k].z); acc[k].w = fmaf(a, cur[u].w, acc[k].w);
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) cur[u] = nxt[u];
  }
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    acc[k].x += __shfl_xor_sync(0xffffffffu, acc[k].x, 16);
    acc[k].y += __shfl_xor_sync(0xffffffffu, acc[k].y, 16);
    acc[k].z += __shfl_xor_sync(0xffffffffu, acc[k].z, 16);
    acc[k].w += __shfl_xor_sync(0xffffffffu, acc[k].w, 16);
    if (lane < 16 && k < rpa)
      *reinterpret_cast<float4*>(red + ((size_t)warp * KMAX + k) * CTX_COLS + c4 * 4) = acc[k];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < rpa * CTX_COLS; i += 256) {
    const int k = i / CTX_COLS, c = i - k * CTX_COLS;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[((size_t)w * KMAX + k) * CTX_COLS + c];
    state[(size_t)(art * d.K + k) * d.SW + blockIdx.x * CTX_COLS + c] = s;
  }
}

template <int KMAX>
static void launch_ctx_t(const Dims& d, int B, int rpa, const int* lens, const int* active,
                         const float* enc_states, const float* e, float* attn, float* state, cudaStream_t st) {
  const size_t smem = (size_t)(KMAX * d.T + 8 * KMAX * CTX_COLS) * sizeof(float);
  static PerDeviceOnce once;
  if (smem > 48 * 1024)
    once.ensure(smem, [&] { cudaFuncSetAttribute(attn_context_kernel<KMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); });
  dim3 grid(d.A / CTX_COLS, B);
  attn_context_kernel<KMAX><<<grid, 256, smem, st>>>(d, rpa, lens, active, enc_states, e, attn, state);
}

int launch_attn_context(const Dims& d, int B, int rpa, const int* lens, const int* active,
                        const float* enc_states, const float* e, float* attn, float* state, cudaStream_t st) {
  if (d.K <= 4) launch_ctx_t<4>(d, B, rpa, lens, active, enc_states, e, attn, state, st);
  else if (d.K <= 8) launch_ctx_t<8>(d, B, rpa, lens, active, enc_states, e, attn, state, st);
  else launch_ctx_t<16>(d, B, rpa, lens, active, enc_states, e, attn, state, st);
  return 1;
}

// ======================================================================================
// p_gen, softmax statistics
// ======================================================================================
__global__ void __launch_bounds__(256)
pgen_kernel(int SW, int K, int R, const float* __restrict__ state, const float* __restrict__ w,
            const float* __restrict__ bias, float* __restrict__ pgen, const int* __restrict__ active) {
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (r >= R || !active[r / K]) return;
  const float* s = state + (size_t)r * SW;
  float acc = 0.f;
  if ((SW & 3) == 0 && SW <= 12 * 128) {            // 128-bit loads, all of a lane's chunks requested before the first FMA
    float4 sv[12], wv[12];
#pragma unroll
    for (int q = 0; q < 12; ++q) {
      const int i = lane + 32 * q;
      const bool in = 4 * i < SW;
      sv[q] = in ? *(reinterpret_cast<const float4*>(s) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      wv[q] = in ? __ldg(reinterpret_cast<const float4*>(w) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int q = 0; q < 12; ++q) {
      acc = fmaf(sv[q].x, wv[q].x, acc); acc = fmaf(sv[q].y, wv[q].y, acc);
      acc = fmaf(sv[q].z, wv[q].z, acc); acc = fmaf(sv[q].w, wv[q].w, acc);
    }
  } else {
    for (int i = lane; i < SW; i += 32) acc = fmaf(s[i], w[i], acc);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) pgen[r] = sigmoidf_(acc + bias[0]);
}

int launch_pgen(const Dims& d, int R, const float* state, const float* pg_w, const float* pg_b, float* pgen,
                const int* active, cudaStream_t st) {
  pgen_kernel<<<(R + 7) / 8, 256, 0, st>>>(d.SW, d.K, R, state, pg_w, pg_b, pgen, active);
  return 1;
}

__device__ __forceinline__ float block_reduce_max(float v, float* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float r = sh[0];
  for (int i = 1; i < nw; ++i) r = fmaxf(r, sh[i]);
  return r;
}
__device__ __forceinline__ float block_reduce_sum(float v, float* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float r = 0.f;
  for (int i = 0; i < nw; ++i) r += sh[i];
  return r;
}

__global__ void __launch_bounds__(256)
softmax_stats_kernel(int V, int K, const float* __restrict__ logits, float* __restrict__ rowmax,
                     float* __restrict__ rowsum, const int* __restrict__ active) {
  __shared__ float sh[8];
  const int r = blockIdx.x;
  if (!active[r / K]) return;
  const float* l = logits + (size_t)r * V;
  float m = -FLT_MAX;
  for (int i = threadIdx.x; i < V; i += 256) m = fmaxf(m, l[i]);
  m = block_reduce_max(m, sh);
  float s = 0.f;
  for (int i = threadIdx.x; i < V; i += 256) s += expf(l[i] - m);
  s = block_reduce_sum(s, sh);
  if (threadIdx.x == 0) { rowmax[r] = m; rowsum[r] = s; }
}

int launch_softmax_stats(const Dims& d, int R, const float* logits, float* rowmax, float* rowsum,
                         const int* active, cudaStream_t st) {
  softmax_stats_kernel<<<R, 256, 0, st>>>(d.V, d.K, logits, rowmax, rowsum, active);
  return 1;
}

// ======================================================================================
// Final distribution + top-2K
// ======================================================================================
__device__ __forceinline__ bool better(float v1, int i1, float v2, int i2) {
  return v1 > v2 || (v1 == v2 && i1 < i2);          // tf.nn.top_k: descending, ties -> lower index
}

template <int K2MAX>
__device__ __forceinline__ void topk_insert(float (&tv)[K2MAX], int (&ti)[K2MAX], int k2, float v, int i) {
  if (!better(v, i, tv[K2MAX - 1], ti[K2MAX - 1])) return;
  tv[K2MAX - 1] = v; ti[K2MAX - 1] = i;
#pragma unroll
  for (int j = K2MAX - 1; j > 0; --j) {
    if (better(tv[j], ti[j], tv[j - 1], ti[j - 1])) {
      const float fv = tv[j]; tv[j] = tv[j - 1]; tv[j - 1] = fv;
      const int fi = ti[j]; ti[j] = ti[j - 1]; ti[j - 1] = fi;
    }
  }
  (void)k2;
}

template <int K2MAX>
__global__ void __launch_bounds__(256)
final_topk_kernel(Dims d, FinalArgs a) {
  extern __shared__ __align__(16) unsigned char smraw[];
  const int r = blockIdx.x, art = r / d.K;
  if (!a.active[art]) return;
  const int V = d.V, T = d.T, K2 = d.K2;
  const int nbm = (V + 31) / 32;
  unsigned* bitmap = reinterpret_cast<unsigned*>(smraw);                  // [nbm]
  float* cv = reinterpret_cast<float*>(bitmap + nbm);                     // [256*K2MAX]
  int* ci = reinterpret_cast<int*>(cv + 256 * K2MAX);                     // [256*K2MAX]
  float* redv = reinterpret_cast<float*>(ci + 256 * K2MAX);               // [8]
  int* redi = reinterpret_cast<int*>(redv + 8);                           // [8]
  int* redw = redi + 8;                                                   // [8] owner thread
  const float* l = a.logits + (size_t)r * V;
  const float m = a.rowmax[r], s = a.rowsum[r];
  const float pg = d.pointer_gen ? a.pgen[r] : 1.f;
  const int nu = d.pointer_gen ? a.uniq_cnt[art] : 0;
  const int* uid = a.uniq_ids + (size_t)art * T;
  for (int i = threadIdx.x; i < nbm; i += 256) bitmap[i] = 0u;
  __syncthreads();
  for (int u = threadIdx.x; u < nu; u += 256) {
    const int id = uid[u];
    if (id < V) atomicOr(&bitmap[id >> 5], 1u << (id & 31));
  }
  __syncthreads();
  float tv[K2MAX];
  int ti[K2MAX];
#pragma unroll
  for (int j = 0; j < K2MAX; ++j) { tv[j] = -FLT_MAX; ti[j] = 0x7fffffff; }
  // vocabulary words that do not occur in the article: p_gen * softmax        (model.py:158)
  for (int w = threadIdx.x; w < V; w += 256) {
    if ((bitmap[w >> 5] >> (w & 31)) & 1u) continue;
    const float val = d.pointer_gen ? pg * (expf(l[w] - m) / s) : expf(l[w] - m) / s;
    topk_insert<K2MAX>(tv, ti, K2, val, w);
  }
  // extended-vocabulary ids that occur in the article: add the copy mass       (model.py:159-181)
  if (nu > 0) {
    const int* ps = a.pos_start + (size_t)art * (T + 1);
    const int* pl = a.pos_list + (size_t)art * T;
    const float* at = a.attn + (size_t)r * T;
    const float omp = 1.0f - pg;
    for (int u = threadIdx.x; u < nu; u += 256) {
      const int id = uid[u];
      float copy = 0.f;
      for (int q = ps[u]; q < ps[u + 1]; ++q) copy += omp * at[pl[q]];     // scatter_nd: duplicates accumulate
      const float vocab = id < V ? pg * (expf(l[id] - m) / s) : 0.f;       // zeros for in-article OOVs (:162-164)
      topk_insert<K2MAX>(tv, ti, K2, vocab + copy, id);
    }
  }
#pragma unroll
  for (int j = 0; j < K2MAX; ++j) { cv[threadIdx.x * K2MAX + j] = tv[j]; ci[threadIdx.x * K2MAX + j] = ti[j]; }
  __syncthreads();
  // K2 rounds of block arg-best over the heads of the 256 sorted lists
  int head = 0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int round = 0; round < K2; ++round) {
    float bv = head < K2MAX ? cv[threadIdx.x * K2MAX + head] : -FLT_MAX;
    int bi = head < K2MAX ? ci[threadIdx.x * K2MAX + head] : 0x7fffffff;
    int bw = threadIdx.x;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      const int ow = __shfl_xor_sync(0xffffffffu, bw, o);
      if (better(ov, oi, bv, bi)) { bv = ov; bi = oi; bw = ow; }
    }
    if (lane == 0) { redv[warp] = bv; redi[warp] = bi; redw[warp] = bw; }
    __syncthreads();
    float fv = redv[0]; int fi = redi[0], fw = redw[0];
#pragma unroll
    for (int w8 = 1; w8 < 8; ++w8)
      if (better(redv[w8], redi[w8], fv, fi)) { fv = redv[w8]; fi = redi[w8]; fw = redw[w8]; }
    if (threadIdx.x == fw) ++head;
    if (threadIdx.x == 0) {
      const bool valid = fi != 0x7fffffff;
      a.topk_ids[(size_t)r * K2 + round] = valid ? fi : 0;
      a.topk_logp[(size_t)r * K2 + round] = valid ? logf(fv) : -INFINITY;    // model.py:285
    }
    __syncthreads();
  }
}

template <int K2MAX>
static void launch_final_t(const Dims& d, int B, const FinalArgs& a, cudaStream_t st) {
  const size_t smem = (size_t)((d.V + 31) / 32) * 4 + (size_t)256 * K2MAX * 8 + 8 * 12 + 16;
  static PerDeviceOnce once;
  if (smem > 48 * 1024)
    once.ensure(smem, [&] { cudaFuncSetAttribute(final_topk_kernel<K2MAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); });
  final_topk_kernel<K2MAX><<<B * d.K, 256, smem, st>>>(d, a);
}

int launch_final_topk(const Dims& d, int B, const FinalArgs& a, cudaStream_t st) {
  if (d.K2 <= 8) launch_final_t<8>(d, B, a, st);
  else if (d.K2 <= 16) launch_final_t<16>(d, B, a, st);
  else launch_final_t<32>(d, B, a, st);
  return 1;
}

// ======================================================================================
// Beam bookkeeping (beam_search.py:100-168)
// ======================================================================================
constexpr int kStop = 3, kStart = 2;
constexpr int BEAM_MAXC = 512;      // K * 2K candidates, K <= 16

__global__ void __launch_bounds__(128)
beam_step_kernel(Dims d, const int* __restrict__ topk_ids, const float* __restrict__ topk_logp, BeamBufs b) {
  __shared__ double keys[BEAM_MAXC];
  __shared__ double sums[BEAM_MAXC];
  __shared__ int order[BEAM_MAXC];
  __shared__ int toks[BEAM_MAXC];
  __shared__ float lps[BEAM_MAXC];
  __shared__ int n_new_s;
  const int art = blockIdx.x;
  if (!b.active[art]) return;
  const int step = b.steps_done[art];                                   // an active article has done every step so far
  const int K = d.K, K2 = d.K2;
  const int num_orig = step == 0 ? 1 : K;                               // :127-129
  const int nc = num_orig * K2;
  for (int c = threadIdx.x; c < nc; c += 128) {
    const int i = c / K2;
    const float lp = topk_logp[(size_t)(art * K + i) * K2 + (c - i * K2)];
    lps[c] = lp;
    toks[c] = topk_ids[(size_t)(art * K + i) * K2 + (c - i * K2)];      // staged: the serial walk below reads shared memory only
    const double s = b.beam_sum[art * K + i] + (double)lp;              // Hypothesis.log_prob (:72-74), float64
    sums[c] = s;
    keys[c] = s / (double)(step + 2);                                   // avg_log_prob (:77-79), [START] counted
  }
  __syncthreads();
  for (int c = threadIdx.x; c < nc; c += 128) {                         // stable descending sort == rank count
    const double kc = keys[c];
    int rank = 0;
    for (int o = 0; o < nc; ++o) rank += (keys[o] > kc || (keys[o] == kc && o < c)) ? 1 : 0;
    order[rank] = c;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int n_new = 0, n_res = b.res_count[art];
    for (int rk = 0; rk < nc; ++rk) {                                   // :145-154
      const int c = order[rk];
      const int i = c / K2;
      const int tok = toks[c];
      const float lp = lps[c];
      if (tok == kStop) {
        if (step >= d.min_dec_steps) {
          const int q = art * K + n_res;
          b.res_sum[q] = sums[c]; b.res_len[q] = step + 2; b.res_step[q] = step; b.res_parent[q] = i;
          b.res_lp[q] = lp;
          ++n_res;
        }
      } else {
        const size_t hq = ((size_t)art * d.S + step) * K + n_new;
        b.hist_tok[hq] = tok; b.hist_par[hq] = i; b.hist_logp[hq] = lp;
        b.sel_parent[art * K + n_new] = i;
        b.sel_token[art * K + n_new] = tok;
        order[n_new] = c;                                               // reuse: slot -> candidate (n_new <= rk)
        ++n_new;
      }
      if (n_new == K || n_res == K) break;
    }
    n_new_s = n_new;
    b.res_count[art] = n_res;
    b.steps_done[art] = step + 1;
    const bool go = (step + 1 < d.S) && (n_res < K);                    // :110
    if (!go) {
      b.active[art] = 0;
      atomicSub(b.n_active, 1);
      if (n_res == 0) { b.best_kind[art] = 1; b.best_idx[art] = 0; }    // :160-162 results = hyps; hyps[0] is best
      else {
        int best = 0;
        double bk = b.res_sum[art * K] / (double)b.res_len[art * K];
        for (int q = 1; q < n_res; ++q) {                               // :165-168 stable: first maximum wins
          const double k2 = b.res_sum[art * K + q] / (double)b.res_len[art * K + q];
          if (k2 > bk) { bk = k2; best = q; }
        }
        b.best_kind[art] = 0; b.best_idx[art] = best;
      }
    }
  }
  __syncthreads();
  for (int sidx = threadIdx.x; sidx < n_new_s; sidx += 128) b.beam_sum[art * K + sidx] = sums[order[sidx]];
}

int launch_beam_step(const Dims& d, int B, const int* topk_ids, const float* topk_logp,
                     const BeamBufs& b, cudaStream_t st) {
  beam_step_kernel<<<B, 128, 0, st>>>(d, topk_ids, topk_logp, b);
  return 1;
}

__global__ void init_beam_kernel(Dims d, int B, const int* __restrict__ lens, BeamBufs b) {
  const int art = blockIdx.x * blockDim.x + threadIdx.x;
  if (art == 0) *b.n_active = 0;
  __threadfence();
  if (art >= B) return;
  const int act = lens[art] > 0 ? 1 : 0;
  b.active[art] = act;
  b.res_count[art] = 0;
  b.steps_done[art] = 0;
  b.best_kind[art] = 2;
  b.best_idx[art] = 0;
  for (int k = 0; k < d.K; ++k) {
    b.sel_parent[art * d.K + k] = 0;
    b.sel_token[art * d.K + k] = kStart;                                // :100
    b.beam_sum[art * d.K + k] = 0.0;                                    // log_probs=[0.0] (:101)
  }
}
__global__ void count_active_kernel(int B, const int* __restrict__ active, int* n_active) {
  int n = 0;
  for (int i = 0; i < B; ++i) n += active[i];
  *n_active = n;
}

int launch_init_beam(const Dims& d, int B, const int* lens, const BeamBufs& b, cudaStream_t st) {
  init_beam_kernel<<<(B + 127) / 128, 128, 0, st>>>(d, B, lens, b);
  count_active_kernel<<<1, 1, 0, st>>>(B, b.active, b.n_active);
  return 2;
}

__global__ void finalize_kernel(Dims d, int B, const int* __restrict__ lens, BeamBufs b, int* __restrict__ out_tokens,
                                int* __restrict__ out_len, float* __restrict__ out_logp, double* __restrict__ out_avg) {
  const int art = blockIdx.x * blockDim.x + threadIdx.x;
  if (art >= B) return;
  const int K = d.K, S = d.S;
  int kind = b.best_kind[art];
  if (lens[art] <= 0) kind = 2;
  int* ot = out_tokens + (size_t)art * S;
  float* ol = out_logp + (size_t)art * S;
  if (kind == 2) {
    out_len[art] = 0; out_avg[art] = 0.0;
    for (int i = 0; i < S; ++i) { ot[i] = 0; ol[i] = 0.f; }
    return;
  }
  int step, slot, n;
  if (kind == 0) {
    const int q = art * K + b.best_idx[art];
    const int rs = b.res_step[q];
    n = rs + 1;
    ot[rs] = kStop;
    ol[rs] = b.res_lp[q];
    step = rs - 1; slot = b.res_parent[q];
    out_avg[art] = b.res_sum[q] / (double)b.res_len[q];
  } else {
    const int sd = b.steps_done[art];
    n = sd; step = sd - 1; slot = 0;
    out_avg[art] = b.beam_sum[art * K] / (double)(sd + 1);
  }
  out_len[art] = n;
  for (int i = n; i < S; ++i) { ot[i] = 0; ol[i] = 0.f; }      // entries past out_len are zero
  while (step >= 0) {
    const size_t hq = ((size_t)art * S + step) * K + slot;
    ot[step] = b.hist_tok[hq];
    ol[step] = b.hist_logp[hq];
    slot = b.hist_par[hq];
    --step;
  }
}

int launch_finalize(const Dims& d, int B, const int* lens, const BeamBufs& b, int* out_tokens, int* out_len,
                    float* out_logp, double* out_avg, cudaStream_t st) {
  finalize_kernel<<<(B + 63) / 64, 64, 0, st>>>(d, B, lens, b, out_tokens, out_len, out_logp, out_avg);
  return 1;
}

__global__ void fill_kernel(float* p, float v, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
int launch_fill(float* p, float v, size_t n, cudaStream_t st) {
  if (n == 0) return 0;
  fill_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p, v, n);
  return 1;
}

}  // namespace pgn

#ifdef PGEN_LIVE_PE
namespace pgn { void live_read_pe(unsigned int* out16, cudaStream_t st) { cudaMemcpyFromSymbolAsync(out16, live_pe, 64, 0, cudaMemcpyDeviceToHost, st); } }
#endif
