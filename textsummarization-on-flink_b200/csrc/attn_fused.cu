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
// persistent grid of one CTA per SM takes the items from a ticket counter (a CTA that started late simply takes
// fewer).  Three kinds of warps per CTA:
//
//   producer (1 warp)  draws the tickets, and streams each item's F tile, h tile, decoder features and coverage
//            into shared memory with cp.async.bulk completing on mbarriers: a ring of three 64 KB tile buffers (one
//            copy per group of four rows; the groups sit 16 B further apart than their size so that "lane = position"
//            reads are conflict-free), as far ahead as "consumed" mbarriers allow;
//   E warps (16)  warp = (pair of beam rows, slice of the 512 features), lane = position: packed FFMA2 over the row
//            pair, tanh(x) = 1 - 2 / (1 + 2^(2 log2e x)) on the special-function unit, four features share one
//            reciprocal (1.25 MUFU operations per tanh) -- this is the part that is bound by the special-function
//            unit, and it runs back to back, item after item;
//   C warps (8)   meanwhile finish the PREVIOUS item: slices' partial sums added in fixed order, softmax inside
//            the segment (its own maximum), p = exp(e - m_seg), ctx_seg[k][a] = sum_t p[t][k] h[t][a] with
//            thread = 2 features, and write (m_seg, sum p, ctx_seg) to a small scratch.
//
// A second, small kernel (one CTA per beam row) then merges the segments in segment order -- M = max m_seg,
// S = sum s_seg e^(m_seg - M) -- and writes attn (softmax, then the reference's re-normalisation) and the context
// vector.  (Merging inside the first kernel, by the CTA that finished an article's last segment, was measured: with
// nearly all of the SM's memory configured as shared there is next to no L1, and the merge's ~60 L2 loads per thread
// took 15k cycles on the C warps, which then held up the tile ring.)
#include <cuda_runtime.h>

#include <cfloat>
#include <cstdint>

#include "kernels.h"

namespace pgn {

namespace {

constexpr int AF_A = 512, AF_SEG = 32;
constexpr int AF_NE = 512, AF_EW = AF_NE / 32;       // E threads / warps
constexpr int AF_CW = 8, AF_NCT = AF_CW * 32;        // C warps / threads (thread = 2 features in the C phase)
constexpr int AF_NT = AF_NE + AF_NCT + 32;           // + one producer warp
constexpr int AF_GRPF = 4 * AF_A + 4;                // pitch of a group of four rows in floats: 8 KB + 16 B
constexpr int AF_TILE = (AF_SEG / 4) * AF_GRPF * 4;  // bytes of one ring buffer
constexpr float kS = 2.885390081777927f;             // 2 log2(e): e^{2x} = 2^{kS x}
constexpr int AF_MAXSEG = 32;                        // max_enc_steps <= 1024
constexpr int AF_MAXB = 1024;                        // articles per device batch this kernel takes (more: the kernel pair)
constexpr int AF_IQ = 4;                             // item slots between the producer and the other warps

template <int K> struct AfCfg {
  static constexpr int NPAIR = K / 2;
  static constexpr int NSLICE = AF_EW / NPAIR;       // warps (feature slices) per row pair
  static constexpr int SLICE = AF_A / NSLICE;        // features per warp
  static constexpr int NDS = K == 4 ? 2 : 1;         // buffers of decoder features + coverage (shared memory is the limit)
  static constexpr int DS_FLOATS = K * AF_A;         // one buffer: [NPAIR][512][2], kS * decoder features
  static constexpr int OFF_PART = 3 * AF_TILE;                       // [NSLICE][K][32] partial energies
  static constexpr int OFF_DS = OFF_PART + NSLICE * K * 32 * 4;      // [NDS][NPAIR][512][2]
  static constexpr int OFF_VS = OFF_DS + NDS * DS_FLOATS * 4;        // [512] -2 v
  static constexpr int OFF_WS = OFF_VS + AF_A * 4;                   // [512] kS w_c
  static constexpr int OFF_CV = OFF_WS + AF_A * 4;                   // [NDS][K][32] coverage of the item's positions
  static constexpr int OFF_P = OFF_CV + NDS * K * 32 * 4;            // [32][K] p = exp(e - m_seg)
  static constexpr int OFF_BAR = OFF_P + 32 * K * 4;                 // mbarriers
  static constexpr int OFF_INT = OFF_BAR + 192;                      // item slots
  static constexpr int OFF_NSG = OFF_INT + 32;          // [B] segments of every live article, [B] positions of its last segment
  static constexpr int SMEM = OFF_NSG + 2 * AF_MAXB;
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
__device__ __forceinline__ void af_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(af_u32(bar)) : "memory");
}
__device__ __forceinline__ void af_expect(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(af_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void af_bulk(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(af_u32(dst)), "l"(src), "r"(bytes), "r"(af_u32(bar)) : "memory");
}

struct AfArgs {
  const int* lens; const int* active;
  const float* enc_feat; const float* enc_states;    // [B, T, 512]
  const float* dpack;                                // [B, K/2, 512, 2] = kS * decoder features, beam rows interleaved in pairs
  const float* cov;                                  // [B*K, T] (nullptr: no coverage term)
  const float* v; const float* w_c;                  // [512]
  float* e;                                          // [B*K, T] energies (scratch)
  float* pctx;                                       // [B, NSEG, K, 512] per-segment context partials (scratch)
  float* segstat;                                    // [B, NSEG, K, 2] per-segment (max, sum) (scratch)
  int* ticket;                                       // work counter: zero before the launch, zero after it
  int B, T, nseg_max;
};

// Position p of a segment lives in row group p >> 2 (four 2 KB rows, contiguous in HBM -> ONE bulk copy per group) and
// the groups are 16 B further apart than their size, so that the eight lanes of a quarter warp -- which hold positions of
// eight different groups, see af_pos() -- read eight different 16-byte bank groups.
__device__ __forceinline__ int af_off(int p) { return (p >> 2) * AF_GRPF + (p & 3) * AF_A; }      // in floats
__device__ __forceinline__ int af_pos(int lane) { return (lane & 7) * 4 + (lane >> 3); }           // E phase: lane -> position

__device__ __forceinline__ void af_sync_e() { asm volatile("bar.sync 1, %0;" ::"n"(AF_NE) : "memory"); }    // E warps only
__device__ __forceinline__ void af_sync_c() { asm volatile("bar.sync 2, %0;" ::"n"(AF_NCT) : "memory"); }   // C warps only

#ifdef PGEN_LIVE_AF
__device__ unsigned int live_af[16];                 // entry / exit counters per warp role (pgen_debug_where)
#define LIVE_AF(i) atomicAdd(&live_af[i], 1u)
#else
#define LIVE_AF(i) ((void)0)
#endif

// Cycle counters for tools/af_dbg.py -- compiled in only with -DPGEN_AF_CYCLES (PGEN_CYCLES=1 bash build.sh); the
// shipped library has none of it.  Slot layout: [CTA][32]: 0..11 E warp 0, 12..21 first C warp, 22..31 producer.
#ifdef PGEN_AF_CYCLES
__device__ long long af_cyc[160 * 32];
#define AF_T(v) const long long v = clock64()
#define AF_ADD(c, t0, t1) c += (t1) - (t0)
#define AF_ONLY(...) __VA_ARGS__
#else
#define AF_T(v)
#define AF_ADD(c, t0, t1)
#define AF_ONLY(...)
#endif

template <int K>
__global__ void __launch_bounds__(AF_NT, 1)
attn_fused_kernel(AfArgs a) {
  using C = AfCfg<K>;
  constexpr int NDS = C::NDS;
  extern __shared__ __align__(128) unsigned char smraw[];
  float* tiles = reinterpret_cast<float*>(smraw);
  float* part = reinterpret_cast<float*>(smraw + C::OFF_PART);
  float* ds = reinterpret_cast<float*>(smraw + C::OFF_DS);
  float* vs = reinterpret_cast<float*>(smraw + C::OFF_VS);
  float* ws = reinterpret_cast<float*>(smraw + C::OFF_WS);
  float* cvs = reinterpret_cast<float*>(smraw + C::OFF_CV);
  float* psm = reinterpret_cast<float*>(smraw + C::OFF_P);
  uint64_t* full = reinterpret_cast<uint64_t*>(smraw + C::OFF_BAR);    // [3] tile landed (transaction bytes)
  uint64_t* empty = full + 3;                                          // [3] tile consumed
  uint64_t* itembar = empty + 3;                                       // [AF_IQ] item slot written
  uint64_t* dsfull = itembar + AF_IQ;                                  // [2] decoder features + coverage landed
  uint64_t* dsfree = dsfull + 2;                                       // [2] ... consumed
  uint64_t* efull = dsfree + 2;                                        // partial energies of an item written
  uint64_t* efree = efull + 1;                                         // ... consumed
  volatile int* items = reinterpret_cast<int*>(smraw + C::OFF_INT);    // [AF_IQ] the CTA's j-th item (-1: no more)
  unsigned char* nsg = smraw + C::OFF_NSG;
  unsigned char* lastn = nsg + AF_MAXB;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int T = a.T, NSEG = a.nseg_max;
  const int n_items = a.B * NSEG;
  AF_T(t_entry);
  AF_ONLY(unsigned long long g_entry; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g_entry));)

  // Segments of every live article (0: finished or empty) and the length of its last segment, in shared memory: with
  // this much of the SM's memory configured as shared there is next to no L1 left, and walking the item list must
  // not cost an L2 round trip per look-up.
  for (int b = tid; b < a.B; b += AF_NT) {
    const int len = a.lens[b];
    nsg[b] = (a.active[b] && len > 0) ? (unsigned char)((len + AF_SEG - 1) / AF_SEG) : 0;
    lastn[b] = (unsigned char)(len > 0 ? ((len - 1) & (AF_SEG - 1)) + 1 : 0);
  }
  if (tid == 0) {
    for (int b = 0; b < 16; ++b) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(af_u32(&full[b])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = tid; i < AF_A; i += AF_NT) { vs[i] = -2.0f * a.v[i]; ws[i] = a.cov ? kS * a.w_c[i] : 0.f; }
  if (!a.cov) for (int i = tid; i < NDS * K * 32; i += AF_NT) cvs[i] = 0.f;
  __syncthreads();
  AF_T(t_setup);
  if (tid == 0) LIVE_AF(0);
  auto item_geom = [&](int i, int& art, int& seg, int& n) {
    art = i / NSEG; seg = i - art * NSEG;
    n = seg == (int)nsg[art] - 1 ? (int)lastn[art] : AF_SEG;
  };
  // the j-th item of this CTA, as handed over by the producer (-1: no more)
  auto get_item = [&](uint32_t j) -> int {
    af_bar_wait(&itembar[j % AF_IQ], (j / AF_IQ) & 1);
    return items[j % AF_IQ];
  };

  // Tile loads are numbered: load 2j = F tile of the CTA's j-th item, 2j + 1 = its h tile; load n uses buffer n % 3.
  if (warp == AF_EW + AF_CW) {
    // ===== producer warp =====
    // Draws the CTA's items from the ticket counter (ticket = article * NSEG + segment; tickets of finished articles
    // or of segments past an article's end are skipped), hands them to the other warps through `items`, and streams
    // their data as far ahead as the buffers allow.
    int* ticket = a.ticket;
    auto draw = [&]() -> int {                                         // lane 0 only
      for (;;) {
        const int t = atomicAdd(ticket, 1);
        if (t >= n_items) {
          // every CTA draws exactly one ticket past the end; the last of them leaves the counter at zero
          if (t == n_items + (int)gridDim.x - 1) *ticket = 0;
          return -1;
        }
        const int art = t / NSEG;
        if (t - art * NSEG < (int)nsg[art]) return t;
      }
    };
    auto hand_over = [&](uint32_t j, int it) {                         // lane 0 only
      items[j % AF_IQ] = it;
      af_arrive(&itembar[j % AF_IQ]);
    };
    auto issue_tile = [&](uint32_t nl, int it, bool states) {
      int art, seg, n;
      item_geom(it, art, seg, n);
      const uint32_t b = nl % 3;
      const float* src = (states ? a.enc_states : a.enc_feat) + ((size_t)art * T + (size_t)seg * AF_SEG) * AF_A;
      float* buf = tiles + (size_t)b * (AF_TILE / 4);
      if (lane == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // the buffer's last readers used ordinary loads
        af_expect(&full[b], (uint32_t)n * AF_A * 4u);
      }
      __syncwarp();
      const int rows = min(4, n - 4 * lane);                           // lane g copies row group g
      if (rows > 0) af_bulk(buf + lane * AF_GRPF, src + (size_t)lane * 4 * AF_A, (uint32_t)rows * AF_A * 4u, &full[b]);
    };
    auto issue_ds = [&](uint32_t j, int it) {                          // decoder features (one copy) + coverage (K copies)
      int art, seg, n;
      item_geom(it, art, seg, n);
      const uint32_t b = j % NDS;
      af_bar_wait(&dsfree[b], ((j / NDS) & 1) ^ 1);
      const uint32_t cvb = a.cov ? (uint32_t)((n + 3) & ~3) * 4u : 0u;
      if (lane == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        af_expect(&dsfull[b], (uint32_t)C::DS_FLOATS * 4u + (uint32_t)K * cvb);
        af_bulk(ds + (size_t)b * C::DS_FLOATS, a.dpack + (size_t)art * C::DS_FLOATS, (uint32_t)C::DS_FLOATS * 4u, &dsfull[b]);
      }
      __syncwarp();
      if (a.cov && lane < K)
        af_bulk(cvs + ((size_t)b * K + lane) * 32, a.cov + (size_t)(art * K + lane) * T + seg * AF_SEG, cvb, &dsfull[b]);
    };
    uint32_t nl = 0, j = 0;                                            // next load, item being loaded
    int cur = 0;
    AF_ONLY(long long p_draw = 0, p_empty = 0, p_ds = 0, p_first = 0;)
    AF_T(pa);
    if (lane == 0) { cur = draw(); hand_over(0, cur); }
    cur = __shfl_sync(0xffffffffu, cur, 0);
    AF_T(pb);
    AF_ADD(p_draw, pa, pb);
    while (cur >= 0) {
      int nxt = 0;
      for (int states = 0; states < 2; ++states, ++nl) {
        AF_T(p0);
        af_bar_wait(&empty[nl % 3], ((nl / 3) & 1) ^ 1);               // the buffer's previous tile has been consumed
        AF_T(p1);
        AF_ADD(p_empty, p0, p1);
        issue_tile(nl, cur, states != 0);
        AF_ONLY(if (nl == 0) p_first = clock64() - t_entry;)
        if (states == 0) {
          AF_T(p2);
          issue_ds(j, cur);
          AF_T(p3);
          AF_ADD(p_ds, p2, p3);
          if (lane == 0) { nxt = draw(); hand_over(j + 1, nxt); }      // the next item, while this one's data flies
          nxt = __shfl_sync(0xffffffffu, nxt, 0);
          AF_T(p4);
          AF_ADD(p_draw, p3, p4);
        }
      }
      cur = nxt;
      ++j;
    }
    AF_ONLY(if (lane == 0) { long long* o = af_cyc + blockIdx.x * 32 + 22; o[0] = p_draw; o[1] = p_empty; o[2] = p_ds; o[3] = p_first;
                             o[4] = clock64() - t_entry; o[5] = j; })
    if (lane == 0) LIVE_AF(1);
    return;
  }

  if (warp < AF_EW) {
    // ===== E warps: partial energies of item after item =====
    const int pair = warp % C::NPAIR, slice = warp / C::NPAIR;
    const int mypos = af_pos(lane);
    const unsigned long long kS2 = af_pk2(kS, kS), one2 = af_pk2(1.0f, 1.0f);
    AF_ONLY(long long e_get = 0, e_ds = 0, e_F = 0, e_cmp = 0, e_free = 0, e_sync = 0, e_first = 0; uint32_t e_items = 0;)
    for (uint32_t j = 0;; ++j) {
      AF_T(e0);
      const int it = get_item(j);
      AF_T(e1);
      AF_ADD(e_get, e0, e1);
      if (it < 0) break;
      const uint32_t db = j % NDS, ld = 2 * j, b = ld % 3;
      af_bar_wait(&dsfull[db], (j / NDS) & 1);
      AF_T(e2);
      AF_ADD(e_ds, e1, e2);
      af_bar_wait(&full[b], (ld / 3) & 1);
      AF_T(e3);
      AF_ADD(e_F, e2, e3);
      AF_ONLY(if (j == 0) e_first = e3 - t_entry; ++e_items;)
      const float* cvp = cvs + ((size_t)db * K + 2 * pair) * 32 + mypos;
      const unsigned long long cv2 = af_pk2(cvp[0], cvp[32]);
      const float* Frow = tiles + (size_t)b * (AF_TILE / 4) + af_off(mypos) + slice * C::SLICE;
      const float4* vs4 = reinterpret_cast<const float4*>(vs + slice * C::SLICE);
      const float4* ws4 = reinterpret_cast<const float4*>(ws + slice * C::SLICE);
      const float4* ds4 = reinterpret_cast<const float4*>(ds + (size_t)db * C::DS_FLOATS + ((size_t)pair * AF_A + slice * C::SLICE) * 2);
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
      AF_T(e4);
      AF_ADD(e_cmp, e3, e4);
      if (j > 0) af_bar_wait(efree, (j - 1) & 1);    // the C warps have read the previous item's partial sums
      AF_T(e5);
      AF_ADD(e_free, e4, e5);
      part[(slice * K + 2 * pair) * 32 + mypos] = s0;
      part[(slice * K + 2 * pair + 1) * 32 + mypos] = s1;
      af_sync_e();                                   // every E warp is done with the F tile, the features and `part` is complete
      if (tid == 0) { af_arrive(&empty[b]); af_arrive(&dsfree[db]); af_arrive(efull); }
      AF_T(e6);
      AF_ADD(e_sync, e5, e6);
    }
    AF_ONLY(if (tid == 0) { long long* o = af_cyc + blockIdx.x * 32; unsigned long long g_exit; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g_exit));
                            o[0] = t_setup - t_entry; o[1] = e_get; o[2] = e_ds; o[3] = e_F; o[4] = e_cmp; o[5] = e_free; o[6] = e_sync;
                            o[7] = e_first; o[8] = e_items; o[9] = clock64() - t_entry; o[10] = (long long)g_entry; o[11] = (long long)g_exit; })
    if (lane == 0) LIVE_AF(2);
    return;
  }

  // ===== C warps: softmax inside the segment, context partial =====
  const int ctid = tid - AF_NE, cw = warp - AF_EW;
  float vsum = 0.f;                                  // sum_a v_a: e = sum v - 2 sum v / (1 + e^{2x})
  for (int i = lane; i < AF_A; i += 32) vsum += vs[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) vsum += __shfl_xor_sync(0xffffffffu, vsum, o);
  vsum *= -0.5f;
  AF_ONLY(long long c_get = 0, c_efull = 0, c_soft = 0, c_H = 0, c_ctx = 0, c_st = 0;)
  for (uint32_t j = 0;; ++j) {
    AF_T(c0);
    const int it = get_item(j);
    AF_T(c1);
    AF_ADD(c_get, c0, c1);
    if (it < 0) break;
    int art, seg, n;
    item_geom(it, art, seg, n);
    const int t0 = seg * AF_SEG;
    // ---------------- energies, softmax inside the segment (lane = position) ----------------
    af_bar_wait(efull, j & 1);
    AF_T(c2);
    AF_ADD(c_efull, c1, c2);
#pragma unroll
    for (int kk = 0; kk < K; kk += AF_CW) {
      const int k = kk + cw;
      if (k >= K) break;                             // (beam 4: the upper four C warps have no row)
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
    af_sync_c();                                     // p complete, `part` read
    if (ctid == 0) af_arrive(efree);
    AF_T(c3);
    AF_ADD(c_soft, c2, c3);
    // ---------------- context partial: thread = 2 features, packed over row pairs ----------------
    {
      const uint32_t ld = 2 * j + 1, b = ld % 3;
      af_bar_wait(&full[b], (ld / 3) & 1);
      AF_T(c4);
      AF_ADD(c_H, c3, c4);
      const float* Hc = tiles + (size_t)b * (AF_TILE / 4) + 2 * ctid;
      unsigned long long acc2[C::NPAIR][2];
#pragma unroll
      for (int p = 0; p < C::NPAIR; ++p) { acc2[p][0] = 0ull; acc2[p][1] = 0ull; }
#pragma unroll 8
      for (int t = 0; t < n; ++t) {
        const float2 h = *reinterpret_cast<const float2*>(Hc + af_off(t));
#pragma unroll
        for (int p2 = 0; p2 < C::NPAIR; p2 += 2) {
          const ulonglong2 pp = *reinterpret_cast<const ulonglong2*>(psm + t * K + 2 * p2);   // rows (2 p2, +1), (2 p2 + 2, +3)
          acc2[p2][0] = af_fma2(pp.x, af_pk2(h.x, h.x), acc2[p2][0]);
          acc2[p2][1] = af_fma2(pp.x, af_pk2(h.y, h.y), acc2[p2][1]);
          acc2[p2 + 1][0] = af_fma2(pp.y, af_pk2(h.x, h.x), acc2[p2 + 1][0]);
          acc2[p2 + 1][1] = af_fma2(pp.y, af_pk2(h.y, h.y), acc2[p2 + 1][1]);
        }
      }
      AF_T(c5);
      AF_ADD(c_ctx, c4, c5);
      float* pc = a.pctx + (size_t)(art * NSEG + seg) * K * AF_A + 2 * ctid;
#pragma unroll
      for (int p = 0; p < C::NPAIR; ++p) {
        float lo0, hi0, lo1, hi1;
        af_upk2(acc2[p][0], lo0, hi0);
        af_upk2(acc2[p][1], lo1, hi1);
        *reinterpret_cast<float2*>(pc + (2 * p) * AF_A) = make_float2(lo0, lo1);
        *reinterpret_cast<float2*>(pc + (2 * p + 1) * AF_A) = make_float2(hi0, hi1);
      }
      af_sync_c();                                   // the h buffer and psm are free
      if (ctid == 0) af_arrive(&empty[b]);
      AF_T(c6);
      AF_ADD(c_st, c5, c6);
    }
  }
  AF_ONLY(if (ctid == 0) { long long* o = af_cyc + blockIdx.x * 32 + 12; o[0] = c_get; o[1] = c_efull; o[2] = c_soft; o[3] = c_H; o[4] = c_ctx;
                           o[5] = c_st; o[6] = clock64() - t_entry; })
  if (lane == 0) LIVE_AF(3);
}

// Second kernel: one CTA per beam row of every live article merges the article's segments (segment order) into the
// row's attn and context vector.  Every load of a thread is requested before its first use.
__global__ void __launch_bounds__(512)
attn_merge_kernel(int K, const int* __restrict__ lens, const int* __restrict__ active, const float* __restrict__ e,
                  const float* __restrict__ pctx, const float* __restrict__ segstat, float* __restrict__ attn,
                  float* __restrict__ ctx, int ctx_ld, int T, int NSEG) {
  __shared__ float wsm[AF_MAXSEG], MS[3], red[16];
  const int r = blockIdx.x, art = r / K, k = r - art * K;
  if (!active[art]) return;
  const int len = lens[art];
  if (len <= 0) return;
  const int nseg = (len + AF_SEG - 1) / AF_SEG;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  constexpr int TPT = 2;                             // positions per thread (T <= 1024)
  float ev[TPT];
#pragma unroll
  for (int q = 0; q < TPT; ++q) {
    const int t = tid + q * 512;
    ev[q] = t < len ? e[(size_t)r * T + t] : 0.f;
  }
  constexpr int SB = 16;                             // segments of context partials in flight per thread
  float pv[SB];
#pragma unroll
  for (int q = 0; q < SB; ++q) pv[q] = pctx[((size_t)(art * NSEG + min(q, nseg - 1)) * K + k) * AF_A + tid];
  if (warp == 0) {                                   // segment statistics -> M, S of the row; lane = segment
    float2 st = make_float2(-FLT_MAX, 0.f);
    if (lane < nseg) st = *reinterpret_cast<const float2*>(segstat + ((size_t)(art * NSEG + lane) * K + k) * 2);
    float M = st.x;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
    const float w = lane < nseg ? expf(st.x - M) : 0.f;
    wsm[lane] = w;
    float S = st.y * w;                              // fixed tree over the segments
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) S += __shfl_xor_sync(0xffffffffu, S, o);
    if (lane == 0) { MS[0] = M; MS[1] = S; }
  }
  __syncthreads();
  const float M = MS[0], S = MS[1];
  // softmax over the article's positions (:98) and its sum for the re-normalisation (:100-101)
  float sum = 0.f;
#pragma unroll
  for (int q = 0; q < TPT; ++q) {
    const int t = tid + q * 512;
    ev[q] = t < len ? expf(ev[q] - M) / S : 0.f;
    sum += ev[q];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) red[warp] = sum;
  // context = sum_seg e^(m_seg - M) ctx_seg / S (segment order), thread = feature
  float cacc = 0.f;
  for (int s0 = 0; s0 < nseg; s0 += SB) {
#pragma unroll
    for (int q = 0; q < SB; ++q)
      if (s0 + q < nseg) cacc = fmaf(wsm[s0 + q], pv[q], cacc);
    if (s0 + SB < nseg) {
#pragma unroll
      for (int q = 0; q < SB; ++q) pv[q] = pctx[((size_t)(art * NSEG + min(s0 + SB + q, nseg - 1)) * K + k) * AF_A + tid];
    }
  }
  __syncthreads();
  if (tid == 0) {
    float s2 = 0.f;
    for (int w = 0; w < 16; ++w) s2 += red[w];
    MS[2] = s2;
  }
  __syncthreads();
  const float s2 = MS[2];
#pragma unroll
  for (int q = 0; q < TPT; ++q) {
    const int t = tid + q * 512;
    if (t < T) attn[(size_t)r * T + t] = t < len ? ev[q] / s2 : 0.f;   // exactly 0 past the article (:99)
  }
  ctx[(size_t)r * ctx_ld + tid] = cacc / S / s2;
}

// dpack[b][p][a][0..1] = kS * dfeat[(b K + 2p + {0,1}), a]: the layout the E warps read, one bulk copy per item
__global__ void __launch_bounds__(256)
attn_pack_dfeat_kernel(int K, int R, const float* __restrict__ dfeat, float* __restrict__ dpack, const int* __restrict__ active) {
  const int i = blockIdx.x * 256 + threadIdx.x;      // (row pair, feature)
  if (i >= (R / 2) * AF_A) return;
  const int rp = i / AF_A, f = i - rp * AF_A;        // rows 2 rp, 2 rp + 1
  if (active && !active[(2 * rp) / K]) return;
  reinterpret_cast<float2*>(dpack)[i] = make_float2(kS * dfeat[(size_t)(2 * rp) * AF_A + f], kS * dfeat[(size_t)(2 * rp + 1) * AF_A + f]);
}

template <int K>
static cudaError_t af_prepare() {
  return cudaFuncSetAttribute(attn_fused_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, AfCfg<K>::SMEM);
}

}  // namespace

bool attn_fused_supported(const Dims& d) {
  return d.A == AF_A && (d.K == 4 || d.K == 8) && d.T <= AF_SEG * AF_MAXSEG && d.T % 4 == 0;   // (16-byte coverage copies)
}
int attn_fused_max_articles() { return AF_MAXB; }

int attn_fused_nseg(const Dims& d) { return (d.T + AF_SEG - 1) / AF_SEG; }

size_t attn_fused_pctx_floats(const Dims& d, int B) { return (size_t)B * attn_fused_nseg(d) * d.K * AF_A; }
size_t attn_fused_segstat_floats(const Dims& d, int B) { return (size_t)B * attn_fused_nseg(d) * d.K * 2; }

cudaError_t attn_fused_prepare(const Dims& d) {
  if (d.K == 4) return af_prepare<4>();
  if (d.K == 8) return af_prepare<8>();
  return cudaErrorInvalidValue;
}

int launch_attn_pack_dfeat(const Dims& d, int B, const float* dfeat, float* dpack, const int* active, cudaStream_t st) {
  const int R = B * d.K, n = (R / 2) * AF_A;
  if (n <= 0) return 0;
  attn_pack_dfeat_kernel<<<(n + 255) / 256, 256, 0, st>>>(d.K, R, dfeat, dpack, active);
  return 1;
}

int launch_attn_fused(const Dims& d, int B, const int* lens, const int* active, const float* enc_feat,
                      const float* enc_states, const float* dpack, const float* cov, const float* v, const float* w_c,
                      float* e, float* pctx, float* segstat, int* ticket, float* attn, float* ctx, int ctx_ld, int n_sms,
                      cudaStream_t st) {
  AfArgs a;
  a.lens = lens; a.active = active; a.enc_feat = enc_feat; a.enc_states = enc_states; a.dpack = dpack;
  a.cov = d.use_cov ? cov : nullptr; a.v = v; a.w_c = w_c; a.e = e; a.pctx = pctx; a.segstat = segstat;
  a.ticket = ticket; a.B = B; a.T = d.T; a.nseg_max = attn_fused_nseg(d);
  const int items = B * a.nseg_max;
  const int grid = items < n_sms ? items : n_sms;
  if (d.K == 4) {
    attn_fused_kernel<4><<<grid, AF_NT, AfCfg<4>::SMEM, st>>>(a);
    attn_merge_kernel<<<B * 4, 512, 0, st>>>(4, lens, active, e, pctx, segstat, attn, ctx, ctx_ld, d.T, a.nseg_max);
  } else {
    attn_fused_kernel<8><<<grid, AF_NT, AfCfg<8>::SMEM, st>>>(a);
    attn_merge_kernel<<<B * 8, 512, 0, st>>>(8, lens, active, e, pctx, segstat, attn, ctx, ctx_ld, d.T, a.nseg_max);
  }
  return 2;
}

}  // namespace pgn

#ifdef PGEN_LIVE_AF
namespace pgn { void live_read_af(unsigned int* out16, cudaStream_t st) { cudaMemcpyFromSymbolAsync(out16, live_af, 64, 0, cudaMemcpyDeviceToHost, st); } }
#endif

#ifdef PGEN_AF_CYCLES
extern "C" __attribute__((visibility("default"))) int pgen_debug_afdbg(void* out) {   // 160 x 32 int64, tools/af_dbg.py
  return (int)cudaMemcpyFromSymbol(out, pgn::af_cyc, sizeof(long long) * 160 * 32);
}
#endif
