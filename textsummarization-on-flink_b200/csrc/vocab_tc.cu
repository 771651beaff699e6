// Vocabulary projection on the 5th-generation tensor cores (tcgen05, kind::tf32) fused with the
// softmax statistics and the candidate selection, plus the merge kernel that re-scores the candidates in
// exact fp32, adds the copy distribution and emits the top-2K (model.py:228-238, 146-183, 280-285).
//
//   scores[r, w] = out[r, :] . w[:, w] + v[w]        r < R = articles x beam,  w < V
//
// The [R, V] logits are never written to HBM.  Per CTA: one 128-row block of `out` stays resident in
// shared memory (TMA, 128B swizzle), vocabulary tiles of 128 words stream through a 4-stage TMA ring,
// one elected thread issues tcgen05.mma into a double-buffered TMEM accumulator, and four epilogue warps
// (thread == row) read it back with tcgen05.ld, add the bias and keep a running (max, sum exp) and a
// sorted top-2K list per row.  Operands are tf32-rounded copies (round to nearest), so the tensor-core
// logits carry ~1e-3 absolute noise: they are used ONLY to pick candidates and for the tail of the
// log-sum-exp.  vocab_merge_kernel recomputes the logits of the candidates and of every word that occurs
// in the article in fp32 from the unrounded operands, and corrects the sum for them.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cfloat>
#include <cstdint>
#include <string>

#include "vocab_tc.h"

namespace pgn {

namespace {

constexpr int BM = 128, BN = 128, BK = 32;          // tile: 128 rows x 128 words x 32 tf32 (= 128 bytes of K)
constexpr int STAGES = 3;
constexpr int TILE_BYTES = 128 * 128;                // one [128 x 128B] operand tile
constexpr int NTHREADS = 320;                        // warp 0: TMA, warp 1: MMA, warps 2..9: epilogue
constexpr int EPI_THREADS = 256;
constexpr float kLog2e = 1.4426950408889634f, kLn2 = 0.6931471805599453f;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 26)) __trap();              // a protocol bug becomes an error, not a hang
  }
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] . B[smem]^T, one 128 x 128 x 8 tf32 MMA
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand tile with 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);           // start address            bits [0,14)
  d |= (uint64_t)1 << 16;                            // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;                  // stride byte offset: 8 rows x 128 B
  d |= (uint64_t)1 << 46;                            // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                            // SWIZZLE_128B
  return d;
}
// instruction descriptor: D fp32, A/B tf32, both K-major, M = 128, N = 128
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

template <int K2MAX>
__device__ __forceinline__ void cand_insert(float (&cv)[K2MAX], int (&ci)[K2MAX], float v, int i) {
  // strictly greater: within a thread ids arrive in increasing order, so ties keep the lower id
  if (!(v > cv[K2MAX - 1])) return;
  cv[K2MAX - 1] = v; ci[K2MAX - 1] = i;
#pragma unroll
  for (int j = K2MAX - 1; j > 0; --j) {
    if (cv[j] > cv[j - 1]) {
      const float fv = cv[j]; cv[j] = cv[j - 1]; cv[j - 1] = fv;
      const int fi = ci[j]; ci[j] = ci[j - 1]; ci[j - 1] = fi;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// GEMM + statistics + candidates.  grid = n_mblocks * n_ranges;  CTA (mb, range) handles the vocabulary
// tiles range, range + n_ranges, ...
// ---------------------------------------------------------------------------------------------
template <int K2MAX, bool STORE>
__global__ void __launch_bounds__(NTHREADS, 1)
vocab_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const float* __restrict__ bias, int R, int V, int KB, int n_tiles, int n_mblocks, int n_ranges,
                float* __restrict__ part_m, float* __restrict__ part_s, float* __restrict__ cand_v,
                int* __restrict__ cand_i, float* __restrict__ logits_out) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  unsigned char* smA = smem;                                   // KB tiles
  unsigned char* smB = smem + (size_t)KB * TILE_BYTES;         // STAGES tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(smB + (size_t)STAGES * TILE_BYTES);
  uint64_t* full = bars;                 // [STAGES]
  uint64_t* empty = bars + STAGES;       // [STAGES]
  uint64_t* a_full = bars + 2 * STAGES;  // [1]
  uint64_t* t_full = a_full + 1;         // [2]
  uint64_t* t_empty = t_full + 2;        // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);
  float* bias_s = reinterpret_cast<float*>(bars + 16);           // [8 epilogue warps][64], 16-byte aligned
  float* xstage = bias_s + 8 * 64;                               // [8 epilogue warps][32 words][32 lanes]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mb = blockIdx.x % n_mblocks, range = blockIdx.x / n_mblocks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(a_full, 1);
    for (int a = 0; a < 2; ++a) { mbar_init(&t_full[a], 1); mbar_init(&t_empty[a], EPI_THREADS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {                                             // TMEM: 2 accumulators x 128 columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      mbar_expect_tx(a_full, (uint32_t)KB * TILE_BYTES);
      for (int kb = 0; kb < KB; ++kb) tma_load_2d(smA + (size_t)kb * TILE_BYTES, &tmA, a_full, kb * BK, mb * BM);
      uint32_t it = 0;
      for (int tile = range; tile < n_tiles; tile += n_ranges) {
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
          mbar_expect_tx(&full[s], TILE_BYTES);
          tma_load_2d(smB + (size_t)s * TILE_BYTES, &tmB, &full[s], kb * BK, tile * BN);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      mbar_wait(a_full, 0);
      uint32_t it = 0, tcount = 0;
      for (int tile = range; tile < n_tiles; tile += n_ranges, ++tcount) {
        const int acc = tcount & 1;
        mbar_wait(&t_empty[acc], ((tcount >> 1) & 1) ^ 1);     // epilogue has drained this accumulator
        tc_fence_after();
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&full[s], (it / STAGES) & 1);
          tc_fence_after();
          const uint64_t ad = smem_desc_sw128(smem_u32(smA + (size_t)kb * TILE_BYTES));
          const uint64_t bd = smem_desc_sw128(smem_u32(smB + (size_t)s * TILE_BYTES));
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)                       // 4 x (K = 8 tf32 = 32 bytes) inside the swizzle atom
            tc_mma_tf32(tmem_base + acc * BN, ad + (uint64_t)(k4 * 2), bd + (uint64_t)(k4 * 2), kIdesc, (kb | k4) != 0);
          tc_commit(&empty[s]);                                // frees the stage when these MMAs retire
        }
        tc_commit(&t_full[acc]);                               // accumulator complete
      }
    }
  } else {
    // ===== epilogue: thread == row, two warps per TMEM lane quarter, each taking 64 of the tile's 128 words.
    // Everything runs in the log2 domain (x2 = logit * log2 e) so that one FFMA applies the bias and one
    // MUFU.EX2 gives the exponential.
    const int ew = warp - 2;                                   // 0..7
    const int q = warp & 3;                                    // TMEM lane quarter this warp may read
    const int half = ew >> 2;                                  // which 64 words of the tile
    const int row = mb * BM + q * 32 + lane;
    float* bsm = bias_s + ew * 64;                             // warp-private bias staging
    float m_run = -FLT_MAX, s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    float cv[K2MAX];
    int ci[K2MAX];
#pragma unroll
    for (int j = 0; j < K2MAX; ++j) { cv[j] = -FLT_MAX; ci[j] = -1; }
    uint32_t tcount = 0;
    for (int tile = range; tile < n_tiles; tile += n_ranges, ++tcount) {
      const int acc = tcount & 1;
      const int nb = tile * BN + half * 64;
      __syncwarp();
      bsm[lane] = (nb + lane < V) ? __ldg(bias + nb + lane) * kLog2e : 0.f;
      bsm[lane + 32] = (nb + 32 + lane < V) ? __ldg(bias + nb + 32 + lane) * kLog2e : 0.f;
      __syncwarp();
      mbar_wait(&t_full[acc], (tcount >> 1) & 1);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + half * 64);
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t raw[32];
        tmem_ld32(tbase + c * 32, raw);
        tmem_ld_wait();
        const int n0 = nb + c * 32;
        float x[32];
        float c0 = -FLT_MAX, c1 = -FLT_MAX, c2 = -FLT_MAX, c3 = -FLT_MAX;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 b4 = *reinterpret_cast<const float4*>(bsm + c * 32 + j);
          x[j + 0] = fmaf(__uint_as_float(raw[j + 0]), kLog2e, b4.x);
          x[j + 1] = fmaf(__uint_as_float(raw[j + 1]), kLog2e, b4.y);
          x[j + 2] = fmaf(__uint_as_float(raw[j + 2]), kLog2e, b4.z);
          x[j + 3] = fmaf(__uint_as_float(raw[j + 3]), kLog2e, b4.w);
          if (n0 + j + 3 >= V) {                               // only the last tile
            if (n0 + j + 0 >= V) x[j + 0] = -FLT_MAX;
            if (n0 + j + 1 >= V) x[j + 1] = -FLT_MAX;
            if (n0 + j + 2 >= V) x[j + 2] = -FLT_MAX;
            if (n0 + j + 3 >= V) x[j + 3] = -FLT_MAX;
          }
          c0 = fmaxf(c0, x[j + 0]); c1 = fmaxf(c1, x[j + 1]); c2 = fmaxf(c2, x[j + 2]); c3 = fmaxf(c3, x[j + 3]);
        }
        const float cm = fmaxf(fmaxf(c0, c1), fmaxf(c2, c3));
        if (STORE) {
          if (row < R)
            for (int j = 0; j < 32; ++j)
              if (n0 + j < V) logits_out[(size_t)row * V + n0 + j] = x[j] * kLn2;
        }
        if (cm > m_run) {
          const float f = exp2f(m_run - cm);
          s0 *= f; s1 *= f; s2 *= f; s3 *= f;
          m_run = cm;
        }
#pragma unroll
        for (int j = 0; j < 32; j += 4) {                      // masked words give exp2(-huge) = 0
          s0 += exp2f(x[j + 0] - m_run); s1 += exp2f(x[j + 1] - m_run);
          s2 += exp2f(x[j + 2] - m_run); s3 += exp2f(x[j + 3] - m_run);
        }
        // Candidates.  Lanes are different rows, so "does this word enter my list" is divergent; to keep the
        // warp's trip count at max-over-lanes (not the union of positions) each lane first builds a bit mask of
        // the words above its threshold, then walks its own set bits.
        float thr = cv[K2MAX - 1];
        if (cm > thr) {
          float* xs = xstage + ew * 1024;
          unsigned pass = 0u;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            xs[j * 32 + lane] = x[j];
            pass |= (x[j] > thr) ? (1u << j) : 0u;
          }
          while (pass) {
            const int j = __ffs(pass) - 1;
            pass &= pass - 1;
            const float v = xs[j * 32 + lane];
            cand_insert<K2MAX>(cv, ci, v, n0 + j);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&t_empty[acc]);
    }
    if (row < R) {
      const size_t pi = (size_t)(range * 2 + half) * R + row;
      part_m[pi] = m_run;
      part_s[pi] = (s0 + s1) + (s2 + s3);
#pragma unroll
      for (int j = 0; j < K2MAX; ++j) { cand_v[pi * K2MAX + j] = cv[j]; cand_i[pi * K2MAX + j] = ci[j]; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * BN));
  }
}

__global__ void round_tf32_kernel(const float* __restrict__ in, float* __restrict__ out, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(in[i]));
  out[i] = __uint_as_float(r);
}

// ---------------------------------------------------------------------------------------------
// Merge: one block per article.
// ---------------------------------------------------------------------------------------------
constexpr int MERGE_EXTRA = 8;      // approx candidates re-scored beyond 2K

__device__ __forceinline__ bool better_vi(float v1, int i1, float v2, int i2) {
  return v1 > v2 || (v1 == v2 && i1 < i2);
}

// exact fp32 logits of every distinct in-vocabulary word of the article, for the K rows of the article
__global__ void __launch_bounds__(256)
vocab_rescore_kernel(MergeArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int art = blockIdx.y;
  if (!a.active[art]) return;
  const int nu = a.uniq_cnt[art];
  const int u0 = blockIdx.x * 32;
  if (u0 >= nu) return;
  const int K = a.K, H = a.H, T = a.T, V = a.V;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* outs = sm;                                             // [K][H]
  for (int i = threadIdx.x; i < K * H; i += 256) outs[i] = a.out[(size_t)art * K * H + i];
  __syncthreads();
  for (int i = 0; i < 4; ++i) {
    const int u = u0 + warp * 4 + i;
    if (u >= nu) break;
    const int id = a.uniq_ids[(size_t)art * T + u];
    if (id >= V) continue;                                      // in-article OOV: no vocabulary term
    const float* wr = a.wT + (size_t)id * H;
    float acc[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) acc[k] = 0.f;
    for (int h = lane * 4; h < H; h += 128) {
      const float4 w4 = __ldg(reinterpret_cast<const float4*>(wr + h));
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        if (k < K) {
          const float4 o4 = *reinterpret_cast<const float4*>(outs + k * H + h);
          acc[k] = fmaf(w4.x, o4.x, acc[k]); acc[k] = fmaf(w4.y, o4.y, acc[k]);
          acc[k] = fmaf(w4.z, o4.z, acc[k]); acc[k] = fmaf(w4.w, o4.w, acc[k]);
        }
      }
    }
    const float b = a.bias[id];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      if (k < K) {
        float v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) a.artl[(size_t)(art * K + k) * T + u] = v + b;
      }
    }
  }
}

constexpr int MERGE_THREADS = 512;
constexpr int MERGE_PPL = 10;       // candidate lists ("parts") per lane in the tournament: up to 320 parts

template <int K2MAX>
__global__ void __launch_bounds__(MERGE_THREADS)
vocab_merge_kernel(MergeArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int art = blockIdx.x;
  if (!a.active[art]) return;
  const int K = a.K, K2 = a.K2, H = a.H, V = a.V, T = a.T, NP = a.n_parts, R = a.R;
  const int C = min(K2 + MERGE_EXTRA, NP * K2MAX);
  const int nu = a.pointer_gen ? a.uniq_cnt[art] : 0;
  const int nc = NP * K2MAX;                                    // approx candidates per row
  constexpr int NW = MERGE_THREADS / 32;
  float* outs = sm;                                             // [K][H]
  float* Ms = outs + K * H;                                     // [K]   max, log2 domain
  float* Ss = Ms + K;                                           // [K]   sum of exp(l - M)
  float* apv = Ss + K;                                          // [K][nc] approx logits (log2 domain), per part descending
  int* api = reinterpret_cast<int*>(apv + (size_t)K * nc);      // [K][nc]
  int* sel_i = api + (size_t)K * nc;                            // [K][C]
  float* sel_a = reinterpret_cast<float*>(sel_i + K * C);       // [K][C] approx logit (log2 domain)
  float* sel_x = sel_a + K * C;                                 // [K][C] exact logit
  float* finv = sel_x + K * C;                                  // [K][C + T]
  int* fini = reinterpret_cast<int*>(finv + (size_t)K * (C + T));   // [K][C + T]
  float* att = reinterpret_cast<float*>(fini + (size_t)K * (C + T));   // [K][T]  attention of the K rows
  float* arl = att + (size_t)K * T;                             // [K][T]  exact logits of the article's words
  int* pss = reinterpret_cast<int*>(arl + (size_t)K * T);       // [T + 1] pos_start
  int* pls = pss + (T + 1);                                     // [T]     pos_list
  int* uids = pls + T;                                          // [T]     uniq_ids
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r0 = art * K;
  const int len = a.lens[art];

  // ---- A: stage everything this article needs in shared memory
  for (int i = threadIdx.x; i < K * H; i += MERGE_THREADS) outs[i] = a.out[(size_t)r0 * H + i];
#pragma unroll 4
  for (int i = threadIdx.x; i < K * nc; i += MERGE_THREADS) {
    const int k = i / nc, c = i - k * nc, pt = c / K2MAX, j = c - pt * K2MAX;
    const size_t pi = ((size_t)pt * R + r0 + k) * K2MAX + j;
    apv[i] = a.cand_v[pi];
    api[i] = a.cand_i[pi];
  }
  if (nu > 0) {
    for (int i = threadIdx.x; i < K * len; i += MERGE_THREADS) {
      const int k = i / len, t = i - k * len;
      att[k * T + t] = a.attn[(size_t)(r0 + k) * T + t];
    }
    for (int i = threadIdx.x; i < K * nu; i += MERGE_THREADS) {
      const int k = i / nu, u = i - k * nu;
      arl[k * T + u] = a.artl[(size_t)(r0 + k) * T + u];
    }
    for (int i = threadIdx.x; i <= nu; i += MERGE_THREADS) pss[i] = a.pos_start[(size_t)art * (T + 1) + i];
    for (int i = threadIdx.x; i < len; i += MERGE_THREADS) pls[i] = a.pos_list[(size_t)art * T + i];
    for (int i = threadIdx.x; i < nu; i += MERGE_THREADS) uids[i] = a.uniq_ids[(size_t)art * T + i];
  }
  for (int k = warp; k < K; k += NW) {                          // combine the per-part (max, sum exp2)
    float M = -FLT_MAX;
    for (int pt = lane; pt < NP; pt += 32) M = fmaxf(M, a.part_m[(size_t)pt * R + r0 + k]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
    float S = 0.f;
    for (int pt = lane; pt < NP; pt += 32)
      S += a.part_s[(size_t)pt * R + r0 + k] * exp2f(a.part_m[(size_t)pt * R + r0 + k] - M);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) S += __shfl_xor_sync(0xffffffffu, S, o);
    if (lane == 0) { Ms[k] = M; Ss[k] = S; }
  }
  __syncthreads();
  // ---- B: top-C approx candidates per row: every part's list is already sorted, so a C-round tournament over
  // the list heads suffices (warp per row; lane owns parts lane, lane+32, ...)
  for (int k = warp; k < K; k += NW) {
    float hv[MERGE_PPL]; int hi[MERGE_PPL], hp[MERGE_PPL];
#pragma unroll
    for (int q = 0; q < MERGE_PPL; ++q) {
      const int pt = lane + 32 * q;
      hp[q] = 0;
      hi[q] = pt < NP ? api[k * nc + pt * K2MAX] : -1;
      hv[q] = pt < NP ? apv[k * nc + pt * K2MAX] : -FLT_MAX;
    }
    for (int round = 0; round < C; ++round) {
      float bv = -FLT_MAX; int bi = 0x7fffffff, bq = -1;
#pragma unroll
      for (int q = 0; q < MERGE_PPL; ++q)
        if (hi[q] >= 0 && better_vi(hv[q], hi[q], bv, bi)) { bv = hv[q]; bi = hi[q]; bq = q; }
      int bl = lane;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        const int ol = __shfl_xor_sync(0xffffffffu, bl, o);
        if (better_vi(ov, oi, bv, bi)) { bv = ov; bi = oi; bl = ol; }
      }
      const bool found = bi != 0x7fffffff;
      if (lane == 0) { sel_i[k * C + round] = found ? bi : -1; sel_a[k * C + round] = bv; }
      if (found && lane == bl) {                                // the winning lane advances that list's head
#pragma unroll
        for (int q = 0; q < MERGE_PPL; ++q) {
          if (q == bq) {
            const int pt = lane + 32 * q;
            ++hp[q];
            if (hp[q] < K2MAX) { hi[q] = api[k * nc + pt * K2MAX + hp[q]]; hv[q] = apv[k * nc + pt * K2MAX + hp[q]]; }
            else hi[q] = -1;
          }
        }
      }
    }
  }
  __syncthreads();
  // ---- C: exact fp32 logits of the K*C (row, candidate) pairs (warp per pair)
  for (int it = warp; it < K * C; it += NW) {
    const int k = it / C, id = sel_i[it];
    float acc = 0.f;
    if (id >= 0) {
      const float* wr = a.wT + (size_t)id * H;
      for (int h = lane * 4; h < H; h += 128) {
        const float4 w4 = __ldg(reinterpret_cast<const float4*>(wr + h));
        const float4 o4 = *reinterpret_cast<const float4*>(outs + k * H + h);
        acc = fmaf(w4.x, o4.x, acc); acc = fmaf(w4.y, o4.y, acc); acc = fmaf(w4.z, o4.z, acc); acc = fmaf(w4.w, o4.w, acc);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) sel_x[it] = id >= 0 ? acc + a.bias[id] : -FLT_MAX;
  }
  __syncthreads();
  // ---- D: exact values replace the tensor-core ones inside the sum of exponentials; M stays the tf32 maximum
  if (threadIdx.x < K) {
    const int k = threadIdx.x;
    float S = Ss[k];
    const float M2 = Ms[k], M = M2 * kLn2;
    for (int c = 0; c < C; ++c)
      if (sel_i[k * C + c] >= 0) S += expf(sel_x[k * C + c] - M) - exp2f(sel_a[k * C + c] - M2);
    Ss[k] = S;
  }
  __syncthreads();
  // ---- E: final distribution on the candidate set (model.py:158-181)
  const int nf = C + nu;
  for (int i = threadIdx.x; i < K * nf; i += MERGE_THREADS) {
    const int k = i / nf, c = i - k * nf;
    const int r = r0 + k;
    const float M = Ms[k] * kLn2, S = Ss[k];
    const float pg = a.pointer_gen ? a.pgen[r] : 1.f;
    float val = -FLT_MAX; int id = -1;
    if (c < C) {
      id = sel_i[k * C + c];
      if (id >= 0) {
        int lo = 0, hi2 = nu;                                   // article words are handled below with their copy mass
        while (lo < hi2) { const int mid = (lo + hi2) >> 1; if (uids[mid] < id) lo = mid + 1; else hi2 = mid; }
        if (lo < nu && uids[lo] == id) id = -1;
        else val = a.pointer_gen ? pg * (expf(sel_x[k * C + c] - M) / S) : expf(sel_x[k * C + c] - M) / S;
      }
    } else {
      const int u = c - C;
      id = uids[u];
      const float omp = 1.0f - pg;
      float copy = 0.f;
      for (int qq = pss[u]; qq < pss[u + 1]; ++qq) copy += omp * att[k * T + pls[qq]];   // scatter_nd: duplicates accumulate in t order
      const float vocab = id < V ? pg * (expf(arl[k * T + u] - M) / S) : 0.f;
      val = vocab + copy;
    }
    finv[i] = val; fini[i] = id;
  }
  __syncthreads();
  // ---- F: top-2K with the tf.nn.top_k tie rule (warp per row)
  for (int k = warp; k < K; k += NW) {
    const int r = r0 + k;
    for (int round = 0; round < K2; ++round) {
      float bv = -FLT_MAX; int bi = 0x7fffffff, bp = -1;
      for (int c = lane; c < nf; c += 32) {
        const int id = fini[k * nf + c];
        if (id >= 0 && better_vi(finv[k * nf + c], id, bv, bi)) { bv = finv[k * nf + c]; bi = id; bp = c; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        const int op = __shfl_xor_sync(0xffffffffu, bp, o);
        if (better_vi(ov, oi, bv, bi)) { bv = ov; bi = oi; bp = op; }
      }
      if (lane == 0) {
        a.topk_ids[(size_t)r * K2 + round] = bp >= 0 ? bi : 0;
        a.topk_logp[(size_t)r * K2 + round] = bp >= 0 ? logf(bv) : -INFINITY;   // model.py:285
        if (bp >= 0) fini[k * nf + bp] = -1;
      }
      __syncwarp();
    }
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

std::string make_map(EncodeTiledFn fn, CUtensorMap* map, const float* base, int rows, int H) {
  cuuint64_t dims[2] = {(cuuint64_t)H, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)H * sizeof(float)};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)BM};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r);
  return "";
}

}  // namespace

size_t VocabTC::smem_bytes() const { return (size_t)(KB + STAGES) * TILE_BYTES + 1024 + 256 + 8 * 64 * 4 + 8 * 1024 * 4; }

std::string VocabTC::init(const float* wT_tf32, int V_, int H_, const float* out_tf32, int Rcap, int K2_, int n_sms) {
  V = V_; H = H_; K2 = K2_;
  if (H % BK != 0 || H > 256) return "tensor-core vocabulary path needs hidden_dim % 32 == 0 and <= 256";
  if (K2 > 16) return "tensor-core vocabulary path needs beam_size <= 8";
  KB = H / BK;
  K2MAX = K2 <= 8 ? 8 : 16;
  n_tiles = (V + BN - 1) / BN;
  sms = n_sms;
  void* fnp = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fnp, cudaEnableDefault, &q) != cudaSuccess || !fnp)
    return "cuTensorMapEncodeTiled is not available from the driver";
  EncodeTiledFn fn = (EncodeTiledFn)fnp;
  std::string err = make_map(fn, &tmA, out_tf32, Rcap, H);
  if (err.empty()) err = make_map(fn, &tmB, wT_tf32, V, H);
  if (!err.empty()) return err;
  cudaError_t e = cudaFuncSetAttribute(vocab_tc_kernel<8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes());
  if (e == cudaSuccess) e = cudaFuncSetAttribute(vocab_tc_kernel<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes());
  if (e == cudaSuccess) e = cudaFuncSetAttribute(vocab_tc_kernel<8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes());
  if (e != cudaSuccess) return std::string("cudaFuncSetAttribute(vocab_tc_kernel): ") + cudaGetErrorString(e);
  ready = true;
  return "";
}

void VocabTC::plan(int R, int* n_mblocks, int* n_ranges) const {
  *n_mblocks = (R + BM - 1) / BM;
  int nr = sms / *n_mblocks;
  // the merge kernel stages every candidate list of an article in shared memory: K rows x 2 parts per range x K2MAX
  // entries x 8 bytes.  Keep that below ~96 KB (only binds for wide beams on small batches).
  const int K = K2 / 2 > 0 ? K2 / 2 : 1;
  const int cap = (96 * 1024) / (16 * K * K2MAX);
  if (nr > cap) nr = cap;
  if (nr < 1) nr = 1;
  if (nr > n_tiles) nr = n_tiles;
  *n_ranges = nr;
}

int VocabTC::launch(int R, const float* bias, float* part_m, float* part_s, float* cand_v, int* cand_i,
                    float* logits_dbg, cudaStream_t st) const {
  int nmb, nr;
  plan(R, &nmb, &nr);
  dim3 grid(nmb * nr);
  if (logits_dbg)
    vocab_tc_kernel<8, true><<<grid, NTHREADS, smem_bytes(), st>>>(tmA, tmB, bias, R, V, KB, n_tiles, nmb, nr, part_m, part_s, cand_v, cand_i, logits_dbg);
  else if (K2MAX == 8)
    vocab_tc_kernel<8, false><<<grid, NTHREADS, smem_bytes(), st>>>(tmA, tmB, bias, R, V, KB, n_tiles, nmb, nr, part_m, part_s, cand_v, cand_i, nullptr);
  else
    vocab_tc_kernel<16, false><<<grid, NTHREADS, smem_bytes(), st>>>(tmA, tmB, bias, R, V, KB, n_tiles, nmb, nr, part_m, part_s, cand_v, cand_i, nullptr);
  return 1;
}

size_t VocabTC::merge_smem(int K, int T, int n_parts) const {
  const int nc = n_parts * K2MAX;
  int C = K2 + MERGE_EXTRA;
  if (C > nc) C = nc;
  size_t fl = (size_t)K * H + 2 * K + 2 * (size_t)K * nc + 3 * (size_t)K * C + 2 * (size_t)K * (C + T) +
              2 * (size_t)K * T + 3 * (size_t)T + 1;
  return fl * 4 + 64;
}

int VocabTC::launch_rescore(const MergeArgs& a, int B, cudaStream_t st) const {
  if (!a.pointer_gen) return 0;
  dim3 grid((a.T + 31) / 32, B);
  vocab_rescore_kernel<<<grid, 256, (size_t)a.K * a.H * sizeof(float), st>>>(a);
  return 1;
}

int VocabTC::launch_merge(MergeArgs a, int B, cudaStream_t st) const {
  a.K2MAX = K2MAX;
  const size_t smem = merge_smem(a.K, a.T, a.n_parts);
  static size_t set8 = 0, set16 = 0;
  if (K2MAX == 8) {
    if (smem > 48 * 1024 && smem > set8) { cudaFuncSetAttribute(vocab_merge_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); set8 = smem; }
    vocab_merge_kernel<8><<<B, MERGE_THREADS, smem, st>>>(a);
  } else {
    if (smem > 48 * 1024 && smem > set16) { cudaFuncSetAttribute(vocab_merge_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); set16 = smem; }
    vocab_merge_kernel<16><<<B, MERGE_THREADS, smem, st>>>(a);
  }
  return 1;
}

int launch_round_tf32(const float* in, float* out, size_t n, cudaStream_t st) {
  if (n == 0) return 0;
  round_tf32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(in, out, n);
  return 1;
}

}  // namespace pgn
