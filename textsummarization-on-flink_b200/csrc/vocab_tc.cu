// Vocabulary projection on the 5th-generation tensor cores (tcgen05, kind::tf32) fused with the
// softmax statistics and the candidate selection, plus the final kernel that re-scores the candidates in
// exact fp32, adds the copy distribution and emits the top-2K (model.py:228-238, 146-183, 280-285).
//
//   scores[r, w] = out[r, :] . w[:, w] + v[w]        r < R = articles x beam,  w < V
//
// The [R, V] logits are never written to HBM.  Per CTA: one 128-row block of `out` stays resident in
// shared memory (TMA, 128B swizzle), vocabulary tiles of 128 words stream through a 4-stage TMA ring,
// one elected thread issues tcgen05.mma into a double-buffered TMEM accumulator, and eight epilogue warps
// (thread == row) read it back with tcgen05.ld, add the bias, keep a running (max, sum exp) per row and
// write the maximum of every group of 8 consecutive words.  Operands are tf32-rounded copies (round to
// nearest), so the tensor-core logits carry ~1e-3 absolute noise: they are used ONLY to pick candidate
// groups and for the tail of the log-sum-exp.  The top 2K elements of a row lie inside its top 2K groups (by
// group maximum), so vocab_final_kernel takes the 2K+8 best groups per row and recomputes all their words
// in fp32 from the unrounded operands (plus the tf32-emulated value the tensor core summed, to swap it out of
// the sum of exponentials), adds the copy distribution and emits the top 2K; vocab_rescore_kernel supplies the
// exact logits of every word that occurs in the article.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cfloat>
#include <algorithm>
#include <cstdint>
#include <string>

#include "kernels.h"
#include "vocab_tc.h"

namespace pgn {

namespace {

constexpr int BM = 128, BN = 256, BK = 32;          // CTA-pair tile: 2 x 128 rows, 256 words (128 staged per CTA), 32 tf32 of K
constexpr int BNH = BN / 2;                          // words of a tile staged by each CTA of the pair
constexpr int STAGES = 4;
constexpr int TILE_BYTES = 128 * 128;                // one [128 rows x 128 B] operand tile (A row block or B half tile)
constexpr int NTHREADS = 320;                        // warp 0: TMA, warp 1: MMA, warps 2..9: epilogue
constexpr int EPI_THREADS = 256;
constexpr float kLog2e = 1.4426950408889634f, kLn2 = 0.6931471805599453f;
constexpr int GW = 8;                                // words per candidate group

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
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
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// TMA load issued by either CTA of a pair into its own shared memory, completing on a barrier that may live in the
// other CTA (a shared::cluster address from mapa)
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint32_t bar_cluster_addr, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ uint32_t mapa_rank(const void* p, uint32_t rank) {     // same offset in CTA `rank` of the cluster
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrives on the same barrier of both CTAs of the pair once every MMA issued so far has retired
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
// D[tmem of both CTAs] (+)= A[smem, 128 rows per CTA] . B[smem, 128 words per CTA]^T: one 256 x 256 x 8 tf32 MMA
__device__ __forceinline__ void tc_mma_tf32_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
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
// instruction descriptor: D fp32, A/B tf32, both K-major, M = 256 (pair), N = 256
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);

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
__device__ __forceinline__ float ex2_ftz(float x) {           // one MUFU.EX2, no denormal fix-up: tiny terms may flush to 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
// packed fp32 pairs (FFMA2 / FADD2): two results per issue slot -- the epilogue is issue-bound
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------
// GEMM + softmax statistics + group maxima.  One cluster = one CTA pair (cta_group::2): CTA `crank` keeps row block
// 2 * mp + crank of `out` resident and stages words [crank * 128, crank * 128 + 128) of every 256-word tile, so
// that each byte of the projection matrix is delivered to exactly one SM and each SM's tensor core reads only
// half of B from its own shared memory.  The leader (rank 0) issues the MMAs for both; both run an epilogue on
// their own 128 rows.  grid = 2 * n_mpairs * n_ranges;  pair (mp, range) handles tiles range, range + n_ranges, ...
// ---------------------------------------------------------------------------------------------
// Entry / exit counters per kernel and warp role, read by pgen_debug_where when a call does not return: compiled in only
// with -DPGEN_LIVE (PGEN_LIVE=1 bash build.sh).
#ifdef PGEN_LIVE_VT
__device__ unsigned int live_vt[16];
#define LIVE_VT(i) atomicAdd(&live_vt[i], 1u)
#else
#define LIVE_VT(i) ((void)0)
#endif

template <bool STORE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
vocab_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const float* __restrict__ bias, int R, int V, int KB, int n_tiles, int n_mpairs, int n_ranges,
                float* __restrict__ part_m, float* __restrict__ part_s, float4* __restrict__ gmax, int gstride,
                float* __restrict__ smax, int sstride, float* __restrict__ logits_out) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  unsigned char* smA = smem;                                   // KB tiles
  unsigned char* smB = smem + (size_t)KB * TILE_BYTES;         // STAGES half tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(smB + (size_t)STAGES * TILE_BYTES);
  uint64_t* full = bars;                 // [STAGES]  leader's: both CTAs' halves have landed
  uint64_t* empty = bars + STAGES;       // [STAGES]  each CTA's: the MMAs reading the stage have retired
  uint64_t* a_full = bars + 2 * STAGES;  // [8]       leader's: k-block kb of both row blocks has landed
  uint64_t* t_full = a_full + 8;         // [2]       each CTA's: accumulator complete
  uint64_t* t_empty = t_full + 2;        // [2]       leader's: both CTAs' epilogues have drained the accumulator
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);
  float* bias_s = reinterpret_cast<float*>(bars + 24);           // [8 epilogue warps][128], 16-byte aligned

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t crank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  const int pair = blockIdx.x >> 1;
  const int mp = pair % n_mpairs, range = pair / n_mpairs;
  const int mb = mp * 2 + (int)crank;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int kb = 0; kb < 8; ++kb) mbar_init(&a_full[kb], 1);
    for (int a = 0; a < 2; ++a) { mbar_init(&t_full[a], 1); mbar_init(&t_empty[a], 2 * (EPI_THREADS / 32)); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {                                             // TMEM: 2 accumulators x 256 columns, in both CTAs
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                          // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) LIVE_VT(0);

  if (warp == 0) {
    // ===== TMA producer (both CTAs; completion is signalled on the leader's barriers) =====
    if (lane == 0) {
      // the row block is loaded k-block by k-block, interleaved with the first tile's stages, so that the first
      // MMA waits for 2 x 16 KB instead of the whole 128 KB block
      uint32_t it = 0;
      for (int tile = range; tile < n_tiles; tile += n_ranges) {
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % STAGES;
          if (it < (uint32_t)KB) {
            if (crank == 0) mbar_expect_tx(&a_full[kb], 2u * TILE_BYTES);
            tma_load_2d_pair(smA + (size_t)kb * TILE_BYTES, &tmA, mapa_rank(&a_full[kb], 0), kb * BK, mb * BM);
          }
          mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
          if (crank == 0) mbar_expect_tx(&full[s], 2u * TILE_BYTES);
          tma_load_2d_pair(smB + (size_t)s * TILE_BYTES, &tmB, mapa_rank(&full[s], 0), kb * BK, tile * BN + (int)crank * BNH);
        }
      }
      LIVE_VT(2);
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader only) =====
    if (lane == 0 && crank == 0) {
      uint32_t it = 0, tcount = 0;
      for (int tile = range; tile < n_tiles; tile += n_ranges, ++tcount) {
        const int acc = tcount & 1;
        mbar_wait(&t_empty[acc], ((tcount >> 1) & 1) ^ 1);     // both epilogues have drained this accumulator
        tc_fence_after();
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % STAGES;
          if (it < (uint32_t)KB) mbar_wait(&a_full[kb], 0);
          mbar_wait(&full[s], (it / STAGES) & 1);
          tc_fence_after();
          const uint64_t ad = smem_desc_sw128(smem_u32(smA + (size_t)kb * TILE_BYTES));
          const uint64_t bd = smem_desc_sw128(smem_u32(smB + (size_t)s * TILE_BYTES));
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)                       // 4 x (K = 8 tf32 = 32 bytes) inside the swizzle atom
            tc_mma_tf32_pair(tmem_base + acc * BN, ad + (uint64_t)(k4 * 2), bd + (uint64_t)(k4 * 2), kIdesc, (kb | k4) != 0);
          tc_commit_pair(&empty[s]);                           // frees the stage in both CTAs when these MMAs retire
        }
        tc_commit_pair(&t_full[acc]);                          // accumulator complete, in both CTAs
      }
      LIVE_VT(3);
    }
  } else {
    // ===== epilogue: thread == row, two warps per TMEM lane quarter, each taking 128 of the tile's 256 words.
    // Everything runs in the log2 domain (x2 = logit * log2 e) so that one FFMA applies the bias and one
    // MUFU.EX2 gives the exponential.
    const int ew = warp - 2;                                   // 0..7
    const int q = warp & 3;                                    // TMEM lane quarter this warp may read
    const int half = ew >> 2;                                  // which 128 words of the tile
    const int row = mb * BM + q * 32 + lane;
    float* bsm = bias_s + ew * 128;                            // warp-private bias staging
    const uint32_t t_empty_l0 = mapa_rank(&t_empty[0], 0), t_empty_l1 = mapa_rank(&t_empty[1], 0);
    const uint64_t kL2 = pack2(kLog2e, kLog2e);
    uint32_t tcount = 0;
    for (int tile = range; tile < n_tiles; tile += n_ranges, ++tcount) {
      const int acc = tcount & 1;
      // (max, sum exp2) are kept per HALF TILE, not per CTA: the partition of the vocabulary over CTAs depends on
      // the number of rows in the batch, and a row's softmax denominator must not
      float m_run = -FLT_MAX;
      uint64_t s01 = pack2(0.f, 0.f), s23 = s01;               // four partial sums of exp2(x - m_run)
      const int nb = tile * BN + half * 128;
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i)
        bsm[lane + 32 * i] = (nb + 32 * i + lane < V) ? __ldg(bias + nb + 32 * i + lane) * kLog2e : 0.f;
      __syncwarp();
      mbar_wait(&t_full[acc], (tcount >> 1) & 1);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + half * 128);
      uint32_t raw[2][32];
      float tmx = -FLT_MAX;                                    // maximum of this thread's 128 words of the tile
      tmem_ld32(tbase, raw[0]);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        tmem_ld_wait();
        if (c < 3) tmem_ld32(tbase + 32 * (c + 1), raw[(c + 1) & 1]);      // in flight while chunk c is processed
        const int n0 = nb + c * 32;
        float x[32];
        float g[4];                                            // maxima of the 4 groups of 8 words of this chunk
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 b4 = *reinterpret_cast<const float4*>(bsm + c * 32 + j);
          unpack2(fma2(pack2(__uint_as_float(raw[c & 1][j + 0]), __uint_as_float(raw[c & 1][j + 1])), kL2, pack2(b4.x, b4.y)), x[j + 0], x[j + 1]);
          unpack2(fma2(pack2(__uint_as_float(raw[c & 1][j + 2]), __uint_as_float(raw[c & 1][j + 3])), kL2, pack2(b4.z, b4.w)), x[j + 2], x[j + 3]);
        }
        if (n0 + 32 > V) {                                     // only in the last tile
#pragma unroll
          for (int j = 0; j < 32; ++j) if (n0 + j >= V) x[j] = -FLT_MAX;
        }
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          g[j >> 3] = fmaxf(fmaxf(fmaxf(x[j], x[j + 1]), fmaxf(x[j + 2], x[j + 3])),
                            fmaxf(fmaxf(x[j + 4], x[j + 5]), fmaxf(x[j + 6], x[j + 7])));
        }
        const float cm = fmaxf(fmaxf(g[0], g[1]), fmaxf(g[2], g[3]));
        if (STORE) {
          if (row < R)
            for (int j = 0; j < 32; ++j)
              if (n0 + j < V) logits_out[(size_t)row * V + n0 + j] = x[j] * kLn2;
        }
        if (row < R && n0 < V)                                 // row-major: the 4 chunks of a tile are 64 contiguous bytes of the row
          gmax[(size_t)row * gstride + (n0 >> 5)] = make_float4(g[0], g[1], g[2], g[3]);
        tmx = fmaxf(tmx, cm);
        if (cm > m_run) {
          const float f = ex2_ftz(m_run - cm);
          const uint64_t f2 = pack2(f, f);
          s01 = mul2(s01, f2); s23 = mul2(s23, f2);
          m_run = cm;
        }
        const uint64_t nm = pack2(-m_run, -m_run);
#pragma unroll
        for (int j = 0; j < 32; j += 4) {                      // masked words give exp2(-huge) = 0
          float y0, y1, y2, y3;
          unpack2(add2(pack2(x[j + 0], x[j + 1]), nm), y0, y1);
          unpack2(add2(pack2(x[j + 2], x[j + 3]), nm), y2, y3);
          s01 = add2(s01, pack2(ex2_ftz(y0), ex2_ftz(y1)));
          s23 = add2(s23, pack2(ex2_ftz(y2), ex2_ftz(y3)));
        }
      }
      if (row < R) {
        smax[(size_t)row * sstride + (nb >> 7)] = tmx;
        const size_t pi = (size_t)(tile * 2 + half) * R + row;
        part_m[pi] = m_run;
        float s0, s1, s2, s3;
        unpack2(s01, s0, s1); unpack2(s23, s2, s3);
        part_s[pi] = (s0 + s1) + (s2 + s3);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(acc ? t_empty_l1 : t_empty_l0);      // one arrival per epilogue warp of either CTA
    }
    if (lane == 0) LIVE_VT(4);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) LIVE_VT(5);
  cluster_sync_all();                                          // neither CTA leaves while the other can still signal it
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * BN));
    if (lane == 0) LIVE_VT(1);
  }
}

// ---------------------------------------------------------------------------------------------
// C[M, N] = A[M, K] . B[N, K]^T on the tensor cores with split operands ("3xTF32"): every fp32 operand is split into
// hi = tf32(x) and lo = tf32(x - hi); a.b ~ a_lo.b_hi + a_hi.b_lo + a_hi.b_hi (the lo.lo term is 2^-22 relative), every
// product exact; the tensor core's fp32 accumulator truncates, which leaves ~1e-5 absolute on O(1) sums over K = 512
// (measured in tests/test_gpu_parity.py::test_features_on_tensor_cores_match_fp32; an fp32 FMA chain gives ~1e-6).  Same CTA-pair structure as the vocabulary kernel (cta_group::2, 256 x 256 x 8
// MMAs, each CTA stages its 128 rows of A and 128 rows of B), but both operands stream: a stage holds
// [A_hi | A_lo | B_hi | B_lo] of one 32-wide k-block, 64 KB.  Work item = (256-row block, 256-column block).
// Used for the once-per-article W_h.h_i features (attention_decoder.py:67).
// ---------------------------------------------------------------------------------------------
constexpr int G3_STAGES = 3;
constexpr int G3_STAGE_BYTES = 4 * TILE_BYTES;

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
gemm_tc3_kernel(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
                const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl,
                int M, int N, int KB, int n_mblk, int n_nblk, float* __restrict__ C, int ldc) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)G3_STAGES * G3_STAGE_BYTES);
  uint64_t* full = bars;                  // [STAGES] leader's: both CTAs' tiles of the stage have landed
  uint64_t* empty = bars + G3_STAGES;     // [STAGES] each CTA's: the MMAs reading the stage have retired
  uint64_t* t_full = bars + 2 * G3_STAGES;   // [2] each CTA's
  uint64_t* t_empty = t_full + 2;            // [2] leader's
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t crank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int n_items = n_mblk * n_nblk;

  if (threadIdx.x == 0) {
    for (int s = 0; s < G3_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&t_full[a], 1); mbar_init(&t_empty[a], 2 * (EPI_THREADS / 32)); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) LIVE_VT(8);

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int item = pair; item < n_items; item += n_pairs) {
        const int mrow = (item / n_nblk) * 2 * BM + (int)crank * BM, nrow = (item % n_nblk) * BN + (int)crank * BNH;
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % G3_STAGES;
          mbar_wait(&empty[s], ((it / G3_STAGES) & 1) ^ 1);
          if (crank == 0) mbar_expect_tx(&full[s], 2u * G3_STAGE_BYTES);
          const uint32_t fl = mapa_rank(&full[s], 0);
          unsigned char* st = smem + (size_t)s * G3_STAGE_BYTES;
          tma_load_2d_pair(st, &tmAh, fl, kb * BK, mrow);
          tma_load_2d_pair(st + TILE_BYTES, &tmAl, fl, kb * BK, mrow);
          tma_load_2d_pair(st + 2 * TILE_BYTES, &tmBh, fl, kb * BK, nrow);
          tma_load_2d_pair(st + 3 * TILE_BYTES, &tmBl, fl, kb * BK, nrow);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && crank == 0) {
      uint32_t it = 0, tcount = 0;
      for (int item = pair; item < n_items; item += n_pairs, ++tcount) {
        const int acc = tcount & 1;
        mbar_wait(&t_empty[acc], ((tcount >> 1) & 1) ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % G3_STAGES;
          mbar_wait(&full[s], (it / G3_STAGES) & 1);
          tc_fence_after();
          const uint32_t sb = smem_u32(smem + (size_t)s * G3_STAGE_BYTES);
          const uint64_t ah = smem_desc_sw128(sb), al = smem_desc_sw128(sb + TILE_BYTES);
          const uint64_t bh = smem_desc_sw128(sb + 2 * TILE_BYTES), bl = smem_desc_sw128(sb + 3 * TILE_BYTES);
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) {                      // small terms first, then the leading product
            const uint64_t o = (uint64_t)(k4 * 2);
            tc_mma_tf32_pair(tmem_base + acc * BN, al + o, bh + o, kIdesc, (kb | k4) != 0);
            tc_mma_tf32_pair(tmem_base + acc * BN, ah + o, bl + o, kIdesc, 1);
            tc_mma_tf32_pair(tmem_base + acc * BN, ah + o, bh + o, kIdesc, 1);
          }
          tc_commit_pair(&empty[s]);
        }
        tc_commit_pair(&t_full[acc]);
      }
    }
  } else {
    // epilogue: thread == row, two warps per TMEM lane quarter, each writing 128 of the tile's 256 columns
    const int ew = warp - 2, q = warp & 3, half = ew >> 2;
    const uint32_t t_empty_l0 = mapa_rank(&t_empty[0], 0), t_empty_l1 = mapa_rank(&t_empty[1], 0);
    uint32_t tcount = 0;
    for (int item = pair; item < n_items; item += n_pairs, ++tcount) {
      const int acc = tcount & 1;
      const int row = (item / n_nblk) * 2 * BM + (int)crank * BM + q * 32 + lane;
      const int col0 = (item % n_nblk) * BN + half * 128;
      mbar_wait(&t_full[acc], (tcount >> 1) & 1);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + half * 128);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t raw[32];
        tmem_ld32(tbase + 32 * c, raw);
        tmem_ld_wait();
        if (row < M) {
          float4* dst = reinterpret_cast<float4*>(C + (size_t)row * ldc + col0 + 32 * c);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            dst[j] = make_float4(__uint_as_float(raw[4 * j]), __uint_as_float(raw[4 * j + 1]), __uint_as_float(raw[4 * j + 2]),
                                 __uint_as_float(raw[4 * j + 3]));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(acc ? t_empty_l1 : t_empty_l0);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) LIVE_VT(10);
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * BN));
    if (lane == 0) LIVE_VT(9);
  }
}

// hi = tf32(x), lo = tf32(x - hi): x = hi + lo up to 2^-22 relative
__global__ void split_tf32_kernel(const float* __restrict__ in, float* __restrict__ hi, float* __restrict__ lo, size_t n) {
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  const float4 x = *reinterpret_cast<const float4*>(in + i);
  float4 h, l;
  h.x = tf32_rna(x.x); h.y = tf32_rna(x.y); h.z = tf32_rna(x.z); h.w = tf32_rna(x.w);
  l.x = tf32_rna(x.x - h.x); l.y = tf32_rna(x.y - h.y); l.z = tf32_rna(x.z - h.z); l.w = tf32_rna(x.w - h.w);
  *reinterpret_cast<float4*>(hi + i) = h;
  *reinterpret_cast<float4*>(lo + i) = l;
}

__global__ void round_tf32_kernel(const float* __restrict__ in, float* __restrict__ out, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(in[i]));
  out[i] = __uint_as_float(r);
}

// Sum of N per-lane values over the warp with N - 1 (+ log2(32 / N)) shuffles instead of 5 N: every stage halves the
// number of values a lane carries.  Afterwards lane l holds the total of value l * N / 32 (order fixed -> deterministic).
template <int N>
__device__ __forceinline__ float warp_multi_reduce(float (&v)[N], int lane) {
  int n = N;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    if (n > 1) {
      const int hlf = n >> 1;
      const bool up = (lane & o) != 0;
#pragma unroll
      for (int i = 0; i < N / 2; ++i) {
        if (i < hlf) {
          const float send = up ? v[i] : v[i + hlf];
          const float keep = up ? v[i + hlf] : v[i];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
      }
      n = hlf;
    } else {
      v[0] += __shfl_xor_sync(0xffffffffu, v[0], o);
    }
  }
  return v[0];
}

// exact fp32 logits of every distinct in-vocabulary word of the article, for the K rows of the article.
// A warp takes 4 words: all of their weight rows are requested before the first is used.
template <int KP>
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
  int ids[4];
  float4 w4[4][2];
  float bb[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int u = u0 + warp * 4 + i;
    const int id = u < nu ? __ldg(a.uniq_ids + (size_t)art * T + u) : -1;
    ids[i] = id < V ? id : -1;                                  // in-article OOV: no vocabulary term
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    bb[i] = ids[i] >= 0 ? __ldg(a.bias + ids[i]) : 0.f;
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int h = lane * 4 + hh * 128;
      w4[i][hh] = (ids[i] >= 0 && h < H) ? __ldg(reinterpret_cast<const float4*>(a.wT + (size_t)ids[i] * H + h))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  float* outs = sm;                                             // [K][H]
  for (int i = threadIdx.x; i < K * H; i += 256) outs[i] = a.out[(size_t)art * K * H + i];
  __syncthreads();
  float acc[4 * KP];
#pragma unroll
  for (int i = 0; i < 4 * KP; ++i) acc[i] = 0.f;
#pragma unroll
  for (int hh = 0; hh < 2; ++hh) {
    const int h = lane * 4 + hh * 128;
    if (h < H) {
#pragma unroll
      for (int k = 0; k < KP; ++k) {
        if (k < K) {
          const float4 o4 = *reinterpret_cast<const float4*>(outs + k * H + h);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float s = acc[i * KP + k];
            s = fmaf(w4[i][hh].x, o4.x, s); s = fmaf(w4[i][hh].y, o4.y, s);
            s = fmaf(w4[i][hh].z, o4.z, s); s = fmaf(w4[i][hh].w, o4.w, s);
            acc[i * KP + k] = s;
          }
        }
      }
    }
  }
  const float tot = warp_multi_reduce<4 * KP>(acc, lane);
  const int idx = lane * (4 * KP) / 32;                         // (word slot, row) this lane ended up with
  const int i = idx / KP, k = idx % KP;
  const bool writer = (4 * KP == 32) || (lane & 1) == 0;
  int id = ids[0]; float b = bb[0];
  if (i == 1) { id = ids[1]; b = bb[1]; } else if (i == 2) { id = ids[2]; b = bb[2]; } else if (i == 3) { id = ids[3]; b = bb[3]; }
  if (writer && id >= 0 && k < K) a.artl[(size_t)(art * K + k) * T + u0 + warp * 4 + i] = tot + b;
}

// ---------------------------------------------------------------------------------------------
// Candidate groups of one decoder row + exact re-scoring of their words.  One block per row.
// ---------------------------------------------------------------------------------------------
constexpr int SEL_THREADS = 512, SEL_WARPS = SEL_THREADS / 32;
constexpr int QCAP = 64;            // groups a row may add to its candidate set when the guard trips (more: dense scan)
static_assert(QCAP * GW == SEL_THREADS, "the dense scan handles one word per thread and chunk");

// Order-preserving map float -> uint32 (larger float <=> larger key) and its inverse.
__device__ __forceinline__ uint32_t f2ord(float x) {
  const uint32_t b = __float_as_uint(x);
  return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o) { return __uint_as_float(o ^ ((o >> 31) ? 0x80000000u : 0xffffffffu)); }

// One warp extracts the `rounds` largest of the 32 keys its lanes hold (0 = nothing) into list[0..rounds), sorted
// descending, with ids[] alongside: one redux.max + one ballot per round, no shuffles.
__device__ __forceinline__ void warp_top_rounds(uint32_t key, int id, int rounds, uint32_t* list, int* ids, int lane) {
  for (int round = 0; round < rounds; ++round) {
    const uint32_t m = __reduce_max_sync(0xffffffffu, key);
    const uint32_t ball = __ballot_sync(0xffffffffu, key == m);
    if (lane == __ffs(ball) - 1) {
      list[round] = m;
      ids[round] = m ? id : -1;
      key = 0u;
    }
  }
}
// Warp 0 merges n_lists (<= 32) descending lists of `len` keys into the `rounds` largest overall: lane l walks list l.
// *last_key receives the key of the last round (0: fewer than `rounds` keys exist).
__device__ __forceinline__ void warp_merge_heads(const uint32_t* lists, const int* ids, int n_lists, int len, int rounds,
                                                 int* out, uint32_t* last_key, int lane) {
  int head = 0;
  uint32_t cur = lane < n_lists ? lists[lane * len] : 0u;
  uint32_t m = 0u;
  for (int round = 0; round < rounds; ++round) {
    m = __reduce_max_sync(0xffffffffu, cur);
    const uint32_t ball = __ballot_sync(0xffffffffu, cur == m);
    if (lane == __ffs(ball) - 1) {
      out[round] = m ? ids[lane * len + head] : -1;
      ++head;
      cur = (m && head < len) ? lists[lane * len + head] : 0u;
    }
  }
  if (lane == 0) *last_key = m;
}

// Exact fp32 logit (and the tf32-emulated one the tensor core summed) of the 8 words of each of the n_groups groups
// listed in glist -> cx / ca / cid [base, base + 8 n_groups).  Warp per word, 4 words in flight per warp.
__device__ __forceinline__ void rescore_groups(const MergeArgs& a, const int* glist, int n_groups, int base,
                                               const float* outs, const float* outt, float* cx, float* ca, int* cid,
                                               int warp, int lane) {
  const int nw = n_groups * GW, H = a.H, V = a.V;
  for (int it0 = warp * 4; it0 < nw; it0 += SEL_WARPS * 4) {
    float4 w4[4][2];
    int ids[4];
    float bb[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int it = it0 + u;
      const int grp = it < nw ? glist[it / GW] : -1;
      const int id = grp >= 0 ? grp * GW + (it % GW) : -1;
      ids[u] = (id >= 0 && id < V) ? id : -1;
      bb[u] = ids[u] >= 0 ? __ldg(a.bias + ids[u]) : 0.f;
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int h = lane * 4 + hh * 128;
        w4[u][hh] = (ids[u] >= 0 && h < H) ? __ldg(reinterpret_cast<const float4*>(a.wT + (size_t)ids[u] * H + h))
                                           : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    float acc[8];                                   // (word u, exact) at 2u, (word u, tf32-emulated) at 2u + 1
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float ex = 0.f, em = 0.f;
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int h = lane * 4 + hh * 128;
        if (h < H) {
          const float4 o4 = *reinterpret_cast<const float4*>(outs + h);
          const float4 t4 = *reinterpret_cast<const float4*>(outt + h);
          const float4 w = w4[u][hh];
          ex = fmaf(w.x, o4.x, ex); ex = fmaf(w.y, o4.y, ex); ex = fmaf(w.z, o4.z, ex); ex = fmaf(w.w, o4.w, ex);
          em = fmaf(tf32_rna(w.x), t4.x, em); em = fmaf(tf32_rna(w.y), t4.y, em);
          em = fmaf(tf32_rna(w.z), t4.z, em); em = fmaf(tf32_rna(w.w), t4.w, em);
        }
      }
      acc[2 * u] = ex; acc[2 * u + 1] = em;
    }
    const float tot = warp_multi_reduce<8>(acc, lane);          // lane l holds value l / 4
    const int vi = lane >> 2, u = vi >> 1, it = it0 + u;
    int id = ids[0]; float b = bb[0];
    if (u == 1) { id = ids[1]; b = bb[1]; } else if (u == 2) { id = ids[2]; b = bb[2]; } else if (u == 3) { id = ids[3]; b = bb[3]; }
    if ((lane & 3) == 0 && it < nw) {
      if (vi & 1) ca[base + it] = id >= 0 ? tot + b : -FLT_MAX;
      else { cx[base + it] = id >= 0 ? tot + b : -FLT_MAX; cid[base + it] = id; }
    }
  }
}

// One block per beam row, after the tensor-core pass:
//   1. the 2K+8 best groups of 8 words by group maximum (tf32 logits), hierarchically and without shuffles: first
//      the best 128-word supergroups (one key per thread, redux.max + ballot rounds per warp, warp 0 merges the
//      warps' sorted lists), then the same over the 16 groups of each selected supergroup -- the best groups
//      necessarily lie inside the best supergroups.  One more round than needed yields the maximum of the best
//      group that was NOT selected;
//   2. exact fp32 logit (and the tf32-emulated one, to take the tensor-core value back out of the sum of
//      exponentials) of the 8 words of every selected group;
//   3. softmax denominator from the per-part (max, sum exp2) of the tensor-core pass, corrected by step 2;
//   4. final distribution on candidates + the article's words (model.py:146-183), top-2K with tf.nn.top_k's tie
//      rule (larger value, then lower id) -> topk_ids / topk_logp (model.py:284-285);
//   5. guard: tf.nn.top_k is exact over V + n_oov (model.py:284), the candidate set came from tf32 maxima.  A word
//      outside the selected groups has a tf32 logit <= u (the best unselected maximum) and therefore an exact logit
//      <= u + eps with eps = 1.1e-3 |out| max_w |w_w| (both operands rounded to 11 significant bits: 2^-10 sum |o w|
//      <= 2^-10 |o| |w|, plus the accumulator's rounding).  If p_gen softmax(u + eps) is below the 2K-th value
//      found, no such word can be in the top 2K and the row is exact as it stands.  Otherwise every unselected
//      group whose maximum could reach that value is re-scored too (at most QCAP of them: "widened") and steps 3-4
//      are repeated; if there are more than QCAP (a flat or tied row), every word of the vocabulary is scored in
//      fp32 by this block ("dense").  Both cases are counted in sel_stats.
enum { GUARD_AUTO = 0, GUARD_WIDEN = 1, GUARD_DENSE = 2, GUARD_OFF = 3 };

__global__ void __launch_bounds__(SEL_THREADS)
vocab_final_kernel(MergeArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int r = blockIdx.x, art = r / a.K;
  if (!a.active[art]) return;
  const int V = a.V, H = a.H, G = a.G, T = a.T, NP = a.n_parts, R = a.R, NC = a.NC, K2 = a.K2;
  const int G1 = G + 1, NCT = NC + QCAP * GW;
  const int nsg = (V + 127) >> 7;                   // supergroups of 128 words (16 groups)
  const int ngroups = (V + GW - 1) / GW;
  const int nu = a.pointer_gen ? a.uniq_cnt[art] : 0;
  const int len = a.lens[art];
  unsigned long long* fin = reinterpret_cast<unsigned long long*>(sm);        // [NCT + T] (value, ~id) keys
  unsigned long long* ck = fin + ((NCT + T + 1) & ~1);   // [16 warps][K2] per-warp winners of the final top-2K
  float* outs = reinterpret_cast<float*>(ck + SEL_WARPS * K2);                // [H] decoder output, exact
  float* outt = outs + H;                           // [H] the tf32-rounded copy the tensor core multiplied
  uint32_t* wk = reinterpret_cast<uint32_t*>(outt + H);                       // [16 warps][G + 1] selection lists: keys
  int* wi = reinterpret_cast<int*>(wk + SEL_WARPS * G1);                      // [16 warps][G + 1]                  ids
  float* pm = reinterpret_cast<float*>(wi + SEL_WARPS * G1);                  // [NP] part maxima
  float* ps = pm + NP;                              // [NP] part sums
  float* att = ps + NP;                             // [T] attention of this row
  float* arl = att + T;                             // [T] exact logits of the article's words
  float* cx = arl + T;                              // [NCT] exact logits of the candidates
  float* ca = cx + NCT;                             // [NCT] tf32-emulated logits of the candidates
  int* cid = reinterpret_cast<int*>(ca + NCT);      // [NCT] candidate word ids
  int* pss = cid + NCT;                             // [T + 1] pos_start
  int* pls = pss + (T + 1);                         // [T] pos_list
  int* uids = pls + T;                              // [T] uniq_ids
  int* sels = uids + T;                             // [G + 1] selected supergroups (+ the best one left out)
  int* sel = sels + G1;                             // [G + 1] selected groups (+ the best one left out)
  int* near = sel + G1;                             // [QCAP] groups added by the guard
  float* MS = reinterpret_cast<float*>(near + QCAP);   // [2] M (natural log domain), S
  float* gf = MS + 2;                               // [2] 2K-th final value of the last pass, |out|
  uint32_t* gi = reinterpret_cast<uint32_t*>(gf + 2);  // [4] best unselected supergroup key, group key, counters
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- 0: everything this row needs, one round of loads
  // (every load of a thread is requested before its first store: one memory round trip, not one per array)
  const int tx = threadIdx.x;
  const float sgv = tx < nsg ? a.smax[(size_t)r * a.sstride + tx] : -FLT_MAX;
  const float v_out = tx < H ? a.out[(size_t)r * H + tx] : 0.f;
  const float v_pm = tx < NP ? a.part_m[(size_t)tx * R + r] : 0.f, v_ps = tx < NP ? a.part_s[(size_t)tx * R + r] : 0.f;
  float v_att = 0.f, v_arl = 0.f;
  int v_pls = 0, v_uid = 0, v_pss = 0;
  if (nu > 0) {
    if (tx < len) { v_att = a.attn[(size_t)r * T + tx]; v_pls = a.pos_list[(size_t)art * T + tx]; }
    if (tx < nu) { v_arl = a.artl[(size_t)r * T + tx]; v_uid = a.uniq_ids[(size_t)art * T + tx]; }
    if (tx <= nu) v_pss = a.pos_start[(size_t)art * (T + 1) + tx];
  }
  if (tx < H) { outs[tx] = v_out; outt[tx] = tf32_rna(v_out); }
  if (tx < NP) { pm[tx] = v_pm; ps[tx] = v_ps; }
  if (nu > 0) {
    if (tx < len) { att[tx] = v_att; pls[tx] = v_pls; }
    if (tx < nu) { arl[tx] = v_arl; uids[tx] = v_uid; }
    if (tx <= nu) pss[tx] = v_pss;
  }
  // sizes beyond one element per thread (max_enc_steps > 511, more than 512 parts)
  for (int h = tx + SEL_THREADS; h < H; h += SEL_THREADS) {
    const float o = a.out[(size_t)r * H + h];
    outs[h] = o;
    outt[h] = tf32_rna(o);
  }
  for (int pt = tx + SEL_THREADS; pt < NP; pt += SEL_THREADS) {
    pm[pt] = a.part_m[(size_t)pt * R + r];
    ps[pt] = a.part_s[(size_t)pt * R + r];
  }
  if (nu > 0) {
    for (int t = tx + SEL_THREADS; t < len; t += SEL_THREADS) {
      att[t] = a.attn[(size_t)r * T + t];
      pls[t] = a.pos_list[(size_t)art * T + t];
    }
    for (int u = tx + SEL_THREADS; u < nu; u += SEL_THREADS) {
      arl[u] = a.artl[(size_t)r * T + u];
      uids[u] = a.uniq_ids[(size_t)art * T + u];
    }
    for (int i = tx + SEL_THREADS; i <= nu; i += SEL_THREADS) pss[i] = a.pos_start[(size_t)art * (T + 1) + i];
  }
  if (threadIdx.x < K2) {                           // defaults, overwritten by the winners in step 4
    a.topk_ids[(size_t)r * K2 + threadIdx.x] = 0;
    a.topk_logp[(size_t)r * K2 + threadIdx.x] = -INFINITY;
  }
  // ---- 1a: the G (+1) best supergroups: per-warp lists, then warp 0 merges them; meanwhile the last warp combines
  // the per-part softmax statistics of the tensor-core pass and the one before it takes the norm of the output
  warp_top_rounds(sgv > -FLT_MAX ? f2ord(sgv) : 0u, threadIdx.x, G1, wk + warp * G1, wi + warp * G1, lane);
  __syncthreads();
  if (warp == 0) warp_merge_heads(wk, wi, SEL_WARPS, G1, G1, sels, &gi[0], lane);
  if (warp == SEL_WARPS - 1) {
    float M2 = -FLT_MAX;
    for (int pt = lane; pt < NP; pt += 32) M2 = fmaxf(M2, pm[pt]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) M2 = fmaxf(M2, __shfl_xor_sync(0xffffffffu, M2, o));
    float S = 0.f;
    for (int pt = lane; pt < NP; pt += 32) S += ps[pt] * exp2f(pm[pt] - M2);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) S += __shfl_xor_sync(0xffffffffu, S, o);
    if (lane == 0) { MS[0] = M2 * kLn2; MS[1] = S; }
  }
  if (warp == SEL_WARPS - 2) {
    float q = 0.f;
    for (int h = lane; h < H; h += 32) q = fmaf(outs[h], outs[h], q);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    if (lane == 0) gf[1] = sqrtf(q);
  }
  __syncthreads();
  // ---- 1b: the G (+1) best groups among the 16 G groups of the G best supergroups (thread = group, 64 contiguous
  // bytes of the row's group maxima per supergroup)
  {
    const int n2 = G * 16, nw2 = (n2 + 31) >> 5;    // <= 512 threads, <= 16 warps
    if (warp < nw2) {
      const int sg = (int)threadIdx.x < n2 ? sels[threadIdx.x >> 4] : -1;
      const int grp = sg >= 0 ? sg * 16 + (int)(threadIdx.x & 15) : -1;
      const float gv = (grp >= 0 && grp < ngroups) ? reinterpret_cast<const float*>(a.gmax + (size_t)r * a.gstride)[grp] : -FLT_MAX;
      warp_top_rounds(gv > -FLT_MAX ? f2ord(gv) : 0u, grp, G1, wk + warp * G1, wi + warp * G1, lane);
    }
    __syncthreads();
    if (warp == 0) warp_merge_heads(wk, wi, nw2, G1, G1, sel, &gi[1], lane);
  }
  __syncthreads();
  // ---- 2: exact fp32 logit and tf32-emulated logit of every word of the selected groups
  rescore_groups(a, sel, G, 0, outs, outt, cx, ca, cid, warp, lane);
  __syncthreads();

  const float M = MS[0];
  const float pg = a.pointer_gen ? a.pgen[r] : 1.f;
  // error bound of a tensor-core logit against the fp32 one (natural-log units), see 5. above
  const float eps = 1.1e-3f * gf[1] * a.wnorm_max + 4e-6f * (fabsf(M) + 1.f);
  int ncand = NC, s_done = 0;
  for (int pass = 0; pass < 2; ++pass) {
    // ---- 3: swap the tensor-core value of every (newly) re-scored word for its exact value inside the sum
    if (warp == 0) {
      float S = 0.f;
      for (int c = s_done + lane; c < ncand; c += 32)
        if (cid[c] >= 0) S += expf(cx[c] - M) - expf(ca[c] - M);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) S += __shfl_xor_sync(0xffffffffu, S, o);
      if (lane == 0) MS[1] += S;
    }
    s_done = ncand;
    if (threadIdx.x < SEL_WARPS * K2) ck[threadIdx.x] = 0ull;    // lists of step 4b: 0 = no entry
    if (threadIdx.x == 0) gf[0] = -1.f;
    __syncthreads();
    // ---- 4a: final distribution on the candidate set (model.py:158-181) as (value, ~id) keys
    const int nf = ncand + nu;
    const float S = MS[1];
    for (int c = threadIdx.x; c < nf; c += SEL_THREADS) {
      float val = -FLT_MAX; int id = -1;
      if (c < ncand) {
        id = cid[c];
        if (id >= 0) {
          int lo = 0, hi2 = nu;                                 // article words are handled below with their copy mass
          while (lo < hi2) { const int mid = (lo + hi2) >> 1; if (uids[mid] < id) lo = mid + 1; else hi2 = mid; }
          if (lo < nu && uids[lo] == id) id = -1;
          else val = a.pointer_gen ? pg * (expf(cx[c] - M) / S) : expf(cx[c] - M) / S;
        }
      } else {
        const int u = c - ncand;
        id = uids[u];
        const float omp = 1.0f - pg;
        float copy = 0.f;
        for (int qq = pss[u]; qq < pss[u + 1]; ++qq) copy += omp * att[pls[qq]];   // scatter_nd: duplicates accumulate in t order
        const float vocab = id < V ? pg * (expf(arl[u] - M) / S) : 0.f;
        val = vocab + copy;
      }
      fin[c] = id >= 0 ? ((unsigned long long)f2ord(val) << 32) | (0xffffffffu - (uint32_t)id) : 0ull;
    }
    __syncthreads();
    // ---- 4b: top-2K by counting: every warp ranks the keys of a contiguous sixteenth among themselves (keys are
    // unique: (value, ~id) with distinct ids) and lists its 2K best in descending order; then every list entry is
    // ranked against the other lists by binary search and the 2K best overall are written out.
    {
      const int cw = (nf + SEL_WARPS - 1) / SEL_WARPS, c0 = warp * cw, c1 = min(nf, c0 + cw);
      for (int cb = c0 + lane; cb < c1; cb += 64) {               // a lane ranks two of its keys per pass
        const unsigned long long k0 = fin[cb], k1 = cb + 32 < c1 ? fin[cb + 32] : 0ull;
        int r0 = 0, r1 = 0;
        for (int c = c0; c < c1; ++c) {
          const unsigned long long o = fin[c];
          r0 += o > k0 ? 1 : 0; r1 += o > k1 ? 1 : 0;
        }
        if (k0 && r0 < K2) ck[warp * K2 + r0] = k0;
        if (k1 && r1 < K2) ck[warp * K2 + r1] = k1;
      }
    }
    __syncthreads();
    if (threadIdx.x < SEL_WARPS * K2) {               // the 2K best of the warps' winners, by counting (keys are unique)
      const unsigned long long mine = ck[threadIdx.x];
      if (mine) {
        int rank = 0;
        const ulonglong2* c2 = reinterpret_cast<const ulonglong2*>(ck);
        for (int j = 0; j < SEL_WARPS * K2 / 2; ++j) {
          const ulonglong2 o = c2[j];
          rank += (o.x > mine ? 1 : 0) + (o.y > mine ? 1 : 0);
        }
        if (rank < K2) {
          const float v = ord2f((uint32_t)(mine >> 32));
          a.topk_ids[(size_t)r * K2 + rank] = (int)(0xffffffffu - (uint32_t)mine);
          a.topk_logp[(size_t)r * K2 + rank] = logf(v);            // model.py:285
          if (rank == K2 - 1) gf[0] = v;
        }
      }
    }
    if (pass == 1 || a.guard == GUARD_OFF) break;
    __syncthreads();
    // ---- 5: guard (every quantity below is block-uniform)
    const uint32_t uk = gi[0] > gi[1] ? gi[0] : gi[1];            // best maximum among the groups left out (0: none)
    const float pstar = gf[0];                                    // -1: fewer than 2K entries so far
    if (uk == 0u && a.guard != GUARD_DENSE) break;                // every group of the row was a candidate
    bool ok = false;
    if (uk != 0u) ok = pstar > pg * (expf(ord2f(uk) * kLn2 + eps - M) / S);
    if (ok && a.guard == GUARD_AUTO) break;
    // smallest exact logit an outside word needs to reach the 2K-th value (with slack for the rounding of log / exp)
    float theta = -FLT_MAX;
    if (pstar > 0.f && pg > 0.f) theta = M + logf(pstar * S / pg) - 1e-5f;
    if (a.guard == GUARD_WIDEN && uk != 0u)                       // test mode: always take in the best group left out
      theta = fminf(theta, ord2f(uk) * kLn2 - 1e-3f);
    bool dense = a.guard == GUARD_DENSE;
    int n_near = 0;
    if (!dense) {
      if (threadIdx.x == 0) gi[2] = 0u;
      __syncthreads();
      const float* grow = reinterpret_cast<const float*>(a.gmax + (size_t)r * a.gstride);
      for (int g = threadIdx.x; g < ngroups; g += SEL_THREADS) {
        const float gv = grow[g];
        if (gv > -FLT_MAX && gv * kLn2 + eps >= theta) {
          bool in = false;
          for (int q = 0; q < G; ++q) in |= sel[q] == g;
          if (!in) {
            const uint32_t idx = atomicAdd(&gi[2], 1u);
            if (idx < (uint32_t)QCAP) near[idx] = g;
          }
        }
      }
      __syncthreads();
      n_near = (int)gi[2];
      if (n_near > QCAP) dense = true;
      else if (n_near == 0) break;
      else {                                                      // ascending group order: the arithmetic below must not
        const int mine = (int)threadIdx.x < n_near ? near[threadIdx.x] : 0;   // depend on who won the atomics
        int rank = 0;
        if ((int)threadIdx.x < n_near) for (int j = 0; j < n_near; ++j) rank += near[j] < mine ? 1 : 0;
        __syncthreads();
        if ((int)threadIdx.x < n_near) near[rank] = mine;
        __syncthreads();
      }
    }
    if (!dense) {
      rescore_groups(a, near, n_near, NC, outs, outt, cx, ca, cid, warp, lane);
      ncand = NC + n_near * GW;
      if (threadIdx.x == 0 && a.sel_stats) atomicAdd(a.sel_stats, 1);
      __syncthreads();
      continue;
    }
    // ---- dense: exact fp32 logit of every word of the vocabulary, QCAP groups (= one word per thread) at a time.
    // Keeps the sum of exponentials (exact for every word) and the 2K best words that do not occur in the article
    // (those are ranked with their copy mass from arl); survivors of a chunk are merged into the list by counting.
    {
      unsigned long long* carry = ck;                             // [2][K2], descending, 0 = empty
      unsigned long long* surv = fin;                             // [<= 512]
      if (threadIdx.x < 2 * K2) carry[threadIdx.x] = 0ull;
      float s_part = 0.f;
      int cb = 0;
      for (int g0 = 0; g0 < ngroups; g0 += QCAP) {
        const int ng = min(QCAP, ngroups - g0);
        if ((int)threadIdx.x < ng) near[threadIdx.x] = g0 + threadIdx.x;
        if (threadIdx.x == 0) gi[3] = 0u;
        if ((int)threadIdx.x < K2) carry[(cb ^ 1) * K2 + threadIdx.x] = 0ull;
        __syncthreads();
        rescore_groups(a, near, ng, NC, outs, outt, cx, ca, cid, warp, lane);
        __syncthreads();
        const unsigned long long thr = carry[cb * K2 + K2 - 1];
        if ((int)threadIdx.x < ng * GW) {
          const int id = cid[NC + threadIdx.x];
          if (id >= 0) {
            const float l = cx[NC + threadIdx.x];
            s_part += expf(l - M);
            int lo = 0, hi2 = nu;
            while (lo < hi2) { const int mid = (lo + hi2) >> 1; if (uids[mid] < id) lo = mid + 1; else hi2 = mid; }
            if (!(lo < nu && uids[lo] == id)) {
              const unsigned long long key = ((unsigned long long)f2ord(l) << 32) | (0xffffffffu - (uint32_t)id);
              if (key > thr) surv[atomicAdd(&gi[3], 1u)] = key;
            }
          }
        }
        __syncthreads();
        const int ns = (int)gi[3];
        if (ns > 0) {
          const int tot = K2 + ns;
          for (int i = threadIdx.x; i < tot; i += SEL_THREADS) {
            const unsigned long long k = i < K2 ? carry[cb * K2 + i] : surv[i - K2];
            if (!k) continue;
            int rank = 0;
            for (int j = 0; j < K2; ++j) rank += carry[cb * K2 + j] > k ? 1 : 0;
            for (int j = 0; j < ns; ++j) rank += surv[j] > k ? 1 : 0;
            if (rank < K2) carry[(cb ^ 1) * K2 + rank] = k;
          }
          cb ^= 1;
        }
        __syncthreads();
      }
      // exact softmax denominator: per-warp sums in lane order, then the 16 warps in order (deterministic)
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s_part += __shfl_xor_sync(0xffffffffu, s_part, o);
      float* wsum = reinterpret_cast<float*>(wk);
      if (lane == 0) wsum[warp] = s_part;
      for (int c = threadIdx.x; c < NCT; c += SEL_THREADS) cid[c] = -1;
      __syncthreads();
      if (threadIdx.x == 0) {
        float Sx = 0.f;
        for (int w = 0; w < SEL_WARPS; ++w) Sx += wsum[w];
        MS[1] = Sx;
        if (a.sel_stats) atomicAdd(a.sel_stats + 1, 1);
      }
      if ((int)threadIdx.x < K2) {                                // the list becomes the candidate set of the second pass
        const unsigned long long k = carry[cb * K2 + threadIdx.x];
        if (k) {
          cid[threadIdx.x] = (int)(0xffffffffu - (uint32_t)k);
          cx[threadIdx.x] = ord2f((uint32_t)(k >> 32));
          ca[threadIdx.x] = cx[threadIdx.x];
        }
      }
      ncand = K2; s_done = K2;                                    // nothing left to swap: S is exact already
      __syncthreads();
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

size_t VocabTC::smem_bytes() const { return (size_t)(KB + STAGES) * TILE_BYTES + 1024 + 256 + 8 * 128 * 4; }

std::string VocabTC::init(const float* wT_tf32, int V_, int H_, const float* out_tf32, int Rcap, int K2_, int n_sms) {
  V = V_; H = H_; K2 = K2_;
  if (H % BK != 0 || H > 256) return "tensor-core vocabulary path needs hidden_dim % 32 == 0 and <= 256";
  if (K2 > 16) return "tensor-core vocabulary path needs beam_size <= 8";
  if (V > SEL_THREADS * 128) return "tensor-core vocabulary path needs vocab_size <= 65536";   // one 128-word supergroup per thread
  KB = H / BK;
  G = K2 + 8;                                          // candidate groups per row: the top 2K lie in the top 2K groups
  n_tiles = (V + BN - 1) / BN;
  sms = n_sms;
  void* fnp = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fnp, cudaEnableDefault, &q) != cudaSuccess || !fnp)
    return "cuTensorMapEncodeTiled is not available from the driver";
  EncodeTiledFn fn = (EncodeTiledFn)fnp;
  std::string err = make_map(fn, &tmA, out_tf32, Rcap, H);
  if (err.empty()) err = make_map(fn, &tmB, wT_tf32, V, H);               // box = one CTA's half tile (128 words)
  if (!err.empty()) return err;
  cudaError_t e = cudaFuncSetAttribute(vocab_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes());
  if (e == cudaSuccess) e = cudaFuncSetAttribute(vocab_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes());
  if (e != cudaSuccess) return std::string("cudaFuncSetAttribute(vocab_tc_kernel): ") + cudaGetErrorString(e);
  ready = true;
  return "";
}

int VocabTC::n_parts() const { return 2 * n_tiles; }

void VocabTC::plan(int R, int* n_mpairs, int* n_ranges) const {
  *n_mpairs = (R + 2 * BM - 1) / (2 * BM);
  int nr = (sms / 2) / *n_mpairs;
  if (nr < 1) nr = 1;
  if (nr > n_tiles) nr = n_tiles;
  // the slowest pair sets the time: with ceil(n_tiles / nr) tiles on the busiest pair, use only as many pairs as
  // that needs (196 tiles: 66 pairs x 3 instead of 74 pairs x 2..3) and leave the other SMs to concurrent batches
  const int per = (n_tiles + nr - 1) / nr;
  nr = (n_tiles + per - 1) / per;
  *n_ranges = nr;
}

int VocabTC::launch(int R, const float* bias, float* part_m, float* part_s, float4* gmax, int gstride,
                    float* smax, int sstride, float* logits_dbg, cudaStream_t st) const {
  int nmp, nr;
  plan(R, &nmp, &nr);
  dim3 grid(2 * nmp * nr);                                     // __cluster_dims__(2, 1, 1): consecutive CTAs are a pair
  if (logits_dbg)
    vocab_tc_kernel<true><<<grid, NTHREADS, smem_bytes(), st>>>(tmA, tmB, bias, R, V, KB, n_tiles, nmp, nr, part_m, part_s, gmax, gstride, smax, sstride, logits_dbg);
  else
    vocab_tc_kernel<false><<<grid, NTHREADS, smem_bytes(), st>>>(tmA, tmB, bias, R, V, KB, n_tiles, nmp, nr, part_m, part_s, gmax, gstride, smax, sstride, nullptr);
  return 1;
}

int VocabTC::launch_rescore(const MergeArgs& a, int B, cudaStream_t st) const {
  if (!a.pointer_gen) return 0;
  dim3 grid((a.T + 31) / 32, B);
  if (a.K <= 4) vocab_rescore_kernel<4><<<grid, 256, (size_t)a.K * a.H * sizeof(float), st>>>(a);
  else vocab_rescore_kernel<8><<<grid, 256, (size_t)a.K * a.H * sizeof(float), st>>>(a);
  return 1;
}

size_t VocabTC::final_smem(int T, int n_parts) const {
  const size_t NCT = (size_t)G * GW + (size_t)QCAP * GW, G1 = (size_t)G + 1;
  return 8 * ((NCT + T + 1) & ~(size_t)1) + 8 * (size_t)SEL_WARPS * K2 + 4 * (2 * (size_t)H + 2 * (size_t)SEL_WARPS * G1 + 2 * (size_t)n_parts +
         2 * (size_t)T + 3 * NCT + (3 * (size_t)T + 1) + 2 * G1 + QCAP + 2 + 2 + 4) + 16;
}

int VocabTC::launch_final(const MergeArgs& a, cudaStream_t st) const {
  const size_t smem = final_smem(a.T, a.n_parts);
  static PerDeviceOnce once;
  if (smem > 48 * 1024)
    once.ensure(smem, [&] { cudaFuncSetAttribute(vocab_final_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); });
  vocab_final_kernel<<<a.R, SEL_THREADS, smem, st>>>(a);
  return 1;
}

// ---- 3xTF32 GEMM
std::string GemmTC3::init(const float* a_hi, const float* a_lo, int M_, const float* b_hi, const float* b_lo, int N_, int K_,
                          int n_sms) {
  M = M_; N = N_; K = K_; sms = n_sms;
  if (K % BK != 0 || N % BN != 0) return "3xTF32 GEMM needs K % 32 == 0 and N % 256 == 0";
  void* fnp = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fnp, cudaEnableDefault, &q) != cudaSuccess || !fnp)
    return "cuTensorMapEncodeTiled is not available from the driver";
  EncodeTiledFn fn = (EncodeTiledFn)fnp;
  std::string err = make_map(fn, &tmAh, a_hi, M, K);
  if (err.empty()) err = make_map(fn, &tmAl, a_lo, M, K);
  if (err.empty()) err = make_map(fn, &tmBh, b_hi, N, K);
  if (err.empty()) err = make_map(fn, &tmBl, b_lo, N, K);
  if (!err.empty()) return err;
  const int smem = G3_STAGES * G3_STAGE_BYTES + 1024 + 256;
  if (cudaFuncSetAttribute(gemm_tc3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
    return "cudaFuncSetAttribute(gemm_tc3_kernel) failed";
  ready = true;
  return "";
}

int GemmTC3::launch(int rows, float* C, int ldc, cudaStream_t st) const {
  const int n_mblk = (rows + 2 * BM - 1) / (2 * BM), n_nblk = N / BN;
  const int pairs = std::max(1, std::min(sms / 2, n_mblk * n_nblk));
  gemm_tc3_kernel<<<2 * pairs, NTHREADS, G3_STAGES * G3_STAGE_BYTES + 1024 + 256, st>>>(tmAh, tmAl, tmBh, tmBl, rows, N, K / BK,
                                                                                    n_mblk, n_nblk, C, ldc);
  return 1;
}

int launch_split_tf32(const float* in, float* hi, float* lo, size_t n, cudaStream_t st) {
  if (n == 0) return 0;
  split_tf32_kernel<<<(unsigned)((n / 4 + 255) / 256), 256, 0, st>>>(in, hi, lo, n);
  return 1;
}

int launch_round_tf32(const float* in, float* out, size_t n, cudaStream_t st) {
  if (n == 0) return 0;
  round_tf32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(in, out, n);
  return 1;
}

}  // namespace pgn

#ifdef PGEN_LIVE_VT
namespace pgn { void live_read_vt(unsigned int* out16, cudaStream_t st) { cudaMemcpyFromSymbolAsync(out16, live_vt, 64, 0, cudaMemcpyDeviceToHost, st); } }
#endif
