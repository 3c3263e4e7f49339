// Fused router, forward (SURVEY.md Appendix A.2-A.4 and the statistics of A.8/A.9) on tcgen05 / TMEM fed by TMA.
//
// The gate product logits[T,E] = x[T,d] . Wg[E,d]^T is the one part of the router that CUDA cores cannot do at HBM
// speed (2*T*d*E FMAs against T*d*2 bytes: for E >= 16 the FMA pipe, and the shared-memory reads of Wg that feed
// it, take longer than streaming x).  So the product runs on the tensor pipe while x streams through a TMA ring
// exactly once:
//   * Wg stays fp32-accurate: a prep kernel splits it into bf16 hi + bf16 lo (w = hi + lo + O(2^-17 |w|)); the MMA
//     issuer accumulates x.hi and x.lo into the same TMEM columns, x is bf16 already, products are exact in fp32.
//   * one CTA = 128-token tiles (persistent over tiles): warp 0 TMA producer, warp 1 MMA issuer, 4*G epilogue warps.
//     The accumulator is [128 tokens][EP experts] fp32 in TMEM, double-buffered so the epilogue of tile i overlaps
//     the loads of tile i+1.
//   * epilogue: G threads per token (the straight-line code of one thread per token did not fit the instruction
//     cache at E = 64 and left one warp per scheduler: ncu v4, 47 us at 16384 tokens) read EP / G logits each from
//     TMEM, add the per-image context bias (A.2), exchange (local max, local sum, local top-k) through shared memory; softmax
//     in fp32, top-k by (logit desc, expert index asc) -- softmax is monotone, so this is the order A.4 asks for
//     without the ties fp32 rounding of p could fabricate -- gate renormalisation, and the per-tile statistics:
//     assignment counts per (slot, expert) and sum_t p[t,e] / sum_t lse^2, reduced in a fixed order (warp shuffle
//     tree, then the four warps in order): deterministic, no float atomics.  The plan kernel scans the tiles.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"

namespace b200moe {

constexpr int kRtTile = kRouterTokPerBlock;  // 128 tokens = one MMA M tile = one router block of the plan
constexpr int kRtXBytes = kRtTile * 64 * 2;  // one k-block of the token tile: [128][64] bf16, 128-byte swizzle
constexpr int kRtSmemBudget = 224 * 1024;
constexpr int kRtMaxImg = 3;                 // images a 128-token tile may touch when the router evaluates the context bias
constexpr int kRtMinTokPerImg = 64;          //   itself: tokens_per_image >= 64 -> at most 3
constexpr int kRtMaxNcd = 256;               //   and n_fields * context_dim <= 256 (embedding rows staged in shared memory)

// G = column groups of the epilogue: a token's EP logits are split over G threads (EP / G columns each), so that the
// epilogue -- the part that bounds this kernel -- runs on 4*G warps per CTA with 1/G of the straight-line code.
template <int EP, int G>
struct RouterCfg {
  static constexpr int kThreads = 64 + 128 * G;             // TMA warp, MMA warp, 4*G epilogue warps
  static constexpr int kWBytes = 2 * EP * 64 * 2;           // hi rows then lo rows, [2*EP][64] bf16
  static constexpr int kStageBytes = kRtXBytes + kWBytes;
  static constexpr int kMaxStages = 6;
  static constexpr int kCtasPerSm = (EP == 64) ? 1 : 2;
  static constexpr int kFixedTail = 2 * 4 * (EP + 1) * 4 /*stat_s*/ + 2 * 8 * 64 * 4 /*cnt_s*/ + 256 /*barriers*/ +
                                    kRtMaxImg * 64 * 4 /*cb_s*/ + kRtMaxImg * kRtMaxNcd * 4 /*emb_s*/;
  static constexpr uint32_t kTmemCols = 2 * EP < 32 ? 32 : 2 * EP;
  // exchange area between the column groups: per token and group the local max, local sum and k candidates (key, index)
  static int xch_bytes(int k) { return G * (2 + 2 * k) * kRtTile * 4; }
  static int stages(int k) {
    const int per_cta = kRtSmemBudget / kCtasPerSm;
    int st = (per_cta - kFixedTail - xch_bytes(k) - 1024) / kStageBytes;
    return st > kMaxStages ? kMaxStages : st;
  }
  static int total(int k) { return stages(k) * kStageBytes + kFixedTail + xch_bytes(k) + 1024; }
};

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr) : "memory");
}
template <int N>
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(N) : "memory"); }

// Column sums over the 32 lanes of a warp of N per-lane values in N + log2(32/N) shuffles instead of 5*N: at each
// step a lane keeps one half of its columns and trades the other half with its partner.  Returns, in every lane, the
// 32-lane sum of column col_of_lane<N>(lane); the tree is fixed, so the result is deterministic.
template <int N>
__device__ __forceinline__ float warp_colsum(float (&v)[N], int lane) {
  int off = 16;
#pragma unroll
  for (int half = N / 2; half >= 1; half >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const float keep = up ? v[i + half] : v[i];
      const float send = up ? v[i] : v[i + half];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
    off >>= 1;
  }
  for (; off >= 1; off >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
  return v[0];
}
template <int N>
__device__ __forceinline__ int col_of_lane(int lane) {   // N = 16: lane bits 4..1; N = 8: lane bits 4..2
  int col = 0, off = 16;
#pragma unroll
  for (int half = N / 2; half >= 1; half >>= 1) {
    if (lane & off) col += half;
    off >>= 1;
  }
  return col;
}

struct RouterArgs {
  CtxArgs ctx;                 // context != null: the bias is evaluated per tile inside the kernel (A.2)
  const float* ctx_bias;
  int tokens_per_image;
  int T, d, E, k, normalize;
  float* logits;
  int32_t* topk_idx;
  float* topk_gate;
  int32_t* block_counts;
  float* block_stats;
  int stages;
};

template <int EP, int G>
__global__ void __launch_bounds__(RouterCfg<EP, G>::kThreads, RouterCfg<EP, G>::kCtasPerSm)
router_fwd_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                  const RouterArgs a) {
  using Cfg = RouterCfg<EP, G>;
  constexpr int kThreads = Cfg::kThreads;
  constexpr int kEpi = 128 * G;
  constexpr int CPT = EP / G;                  // columns (experts) per epilogue thread
  constexpr uint32_t kIdesc = make_idesc_bf16(kRtTile, EP, false, false);
  const int kStages = a.stages;

  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  unsigned char* tail = smem + kStages * Cfg::kStageBytes;
  float* stat_s = reinterpret_cast<float*>(tail);                            // [2][4 warps][EP+1]
  int* cnt_s = reinterpret_cast<int*>(stat_s + 2 * 4 * (EP + 1));           // [2][k*E] (<= 8*64)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(cnt_s + 2 * 8 * 64);
  uint64_t* empty_bar = full_bar + Cfg::kMaxStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kMaxStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  float* cb_s = reinterpret_cast<float*>(tail + Cfg::kFixedTail - kRtMaxImg * (64 + kRtMaxNcd) * 4);   // [kRtMaxImg][64]
  float* emb_s = cb_s + kRtMaxImg * 64;                                                                // [kRtMaxImg][ncd]
  float* xch = reinterpret_cast<float*>(tail + Cfg::kFixedTail);             // [G][2 + 2k][128]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int E = a.E, k = a.k, T = a.T;
  const int kE = k * E;
  const int ntiles = (T + kRtTile - 1) / kRtTile;
  const int kblocks = a.d / 64;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
    for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&tfull_bar[s], 1); mbar_init(&tempty_bar[s], kEpi); }
    fence_mbar_init();
  }
  for (int i = threadIdx.x; i < 2 * 8 * 64; i += kThreads) cnt_s[i] = 0;
  if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_sync();                                   // prologue done; inputs of the previous kernel from here on

  if (warp == 0) {
    // ======================= TMA producer =======================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          unsigned char* sx = smem + stage * Cfg::kStageBytes;
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          tma_load_2d(sx, &tmX, &full_bar[stage], kb * 64, tile * kRtTile);   // rows past T arrive as zeros
          tma_load_2d(sx + kRtXBytes, &tmW, &full_bar[stage], kb * 64, 0);    // hi rows 0..EP-1, lo rows EP..2EP-1
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ======================= MMA issuer =======================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * EP;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sx = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint64_t dx = make_smem_desc(sx, 16, 1024);
          const uint64_t dhi = make_smem_desc(sx + kRtXBytes, 16, 1024);
          const uint64_t dlo = make_smem_desc(sx + kRtXBytes + EP * 128, 16, 1024);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            umma_bf16_ss(tmem_d, dx + 2 * ks, dhi + 2 * ks, kIdesc, (kb | ks) != 0 ? 1u : 0u);
            umma_bf16_ss(tmem_d, dx + 2 * ks, dlo + 2 * ks, kIdesc, 1u);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull_bar[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ======================= epilogue: G threads per token, CPT = EP / G experts each =======================
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may read
    const int grp = (warp - 2) >> 2;              // column group (warp-uniform)
    const int et = threadIdx.x - 64;              // 0..kEpi-1 among the epilogue threads
    const int tok = quarter * 32 + lane;          // token row within the tile
    const int c0 = grp * CPT;
    float* xm = xch + (grp * (2 + 2 * k)) * kRtTile + tok;            // local max; +128: local sum;
    uint32_t* xo = reinterpret_cast<uint32_t*>(xm) + 2 * kRtTile;     // then k candidate keys, then k candidate indices
    const int gstride = (2 + 2 * k) * kRtTile;                        // floats between the groups of one token
    int acc = 0;
    uint32_t acc_phase = 0;
    int buf = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int t = tile * kRtTile + tok;
      const bool live = t < T;
      float cbv[CPT];                             // context bias fetched ahead of the accumulator wait
      if (a.ctx.context != nullptr) {
        // A.2 evaluated here: the (at most kRtMaxImg) images this tile touches x E experts, one warp per dot product,
        // while the tile's tokens are still in flight; same arithmetic as the stand-alone kernel (ctx_bias_dot)
        const int b0 = (tile * kRtTile) / a.tokens_per_image;
        const int t_last = min(T, (tile + 1) * kRtTile) - 1;
        const int nimg = t_last / a.tokens_per_image - b0 + 1;
        const int ncd = a.ctx.n * a.ctx.cd;
        // embedding rows of the tile's images first (id -> table value: two dependent loads, every thread in parallel;
        // a dot product that walked them itself serialised five such chains: +6 us at 3200 tokens)
        for (int q = et; q < nimg * ncd; q += kEpi) {
          const int img = q / ncd;
          emb_s[img * kRtMaxNcd + (q - img * ncd)] = ctx_emb_value(a.ctx, b0 + img, q - img * ncd);
        }
        epi_bar_sync<kEpi>();
        for (int p = warp - 2; p < nimg * E; p += kEpi / 32) {
          const int img = p / E, e = p - img * E;
          const float* em = emb_s + img * kRtMaxNcd;
          const float s = ctx_bias_dot([&](int i) { return em[i]; }, a.ctx.wc + static_cast<size_t>(e) * ncd, ncd, lane);
          if (lane == 0) cb_s[img * 64 + e] = s;
        }
        epi_bar_sync<kEpi>();
        const float* row = cb_s + (live ? t / a.tokens_per_image - b0 : 0) * 64 + c0;
#pragma unroll
        for (int c = 0; c < CPT; ++c) cbv[c] = (live && c0 + c < E) ? row[c] : 0.f;
      } else {
        const float* cb = (a.ctx_bias && live) ? a.ctx_bias + static_cast<size_t>(t / a.tokens_per_image) * E + c0 : nullptr;
#pragma unroll
        for (int c = 0; c < CPT; ++c) cbv[c] = (cb && c0 + c < E) ? __ldg(cb + c) : 0.f;
      }
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * EP + c0;
      uint32_t raw[CPT];
      if (CPT == 16) tmem_ld16(taddr, raw);
      else tmem_ld8(taddr, raw);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&tempty_bar[acc]);              // accumulator is in registers: the MMA warp may reuse it
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }

      // logits (A.2), their local maximum, local exponentials
      float lg[CPT];
      float mloc = -INFINITY;
#pragma unroll
      for (int c = 0; c < CPT; ++c) {
        lg[c] = (c0 + c < E) ? __uint_as_float(raw[c]) + cbv[c] : -INFINITY;
        mloc = fmaxf(mloc, lg[c]);
      }
      if (live) {
        float* lo = a.logits + static_cast<size_t>(t) * E + c0;
        if ((E & 3) == 0) {
#pragma unroll
          for (int c = 0; c < CPT; c += 4)
            if (c0 + c < E) *reinterpret_cast<float4*>(lo + c) = make_float4(lg[c], lg[c + 1], lg[c + 2], lg[c + 3]);
        } else {
#pragma unroll
          for (int c = 0; c < CPT; ++c)
            if (c0 + c < E) lo[c] = lg[c];
        }
      }
      const float mref = (mloc == -INFINITY) ? 0.f : mloc;      // a group of padding columns only
      float ex[CPT];
      float sloc = 0.f;
#pragma unroll
      for (int c = 0; c < CPT; ++c) {
        ex[c] = (c0 + c < E) ? expf(lg[c] - mref) : 0.f;
        sloc += ex[c];
      }
      // local top-k by (logit desc, index asc): order-preserving 32-bit image of the fp32 logit (0 = below everything:
      // padding and entries already taken), a max tree over (image, 63 - column) per slot
      {
        uint32_t ord[CPT];
#pragma unroll
        for (int c = 0; c < CPT; ++c) {
          const uint32_t u = __float_as_uint(lg[c]);
          ord[c] = (c0 + c < E) ? ((u & 0x80000000u) ? ~u : (u | 0x80000000u)) : 0u;
        }
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          if (s < k) {
            unsigned long long m[CPT / 2];
#pragma unroll
            for (int i = 0; i < CPT / 2; ++i) {
              const unsigned long long ka = (static_cast<unsigned long long>(ord[i]) << 32) | static_cast<unsigned>(63 - i);
              const unsigned long long kb =
                  (static_cast<unsigned long long>(ord[i + CPT / 2]) << 32) | static_cast<unsigned>(63 - i - CPT / 2);
              m[i] = ka > kb ? ka : kb;
            }
#pragma unroll
            for (int w = CPT / 4; w >= 1; w >>= 1) {
#pragma unroll
              for (int i = 0; i < w; ++i) m[i] = m[i] > m[i + w] ? m[i] : m[i + w];
            }
            const int bi = 63 - static_cast<int>(m[0] & 63ull);
#pragma unroll
            for (int c = 0; c < CPT; ++c) ord[c] = (c == bi) ? 0u : ord[c];
            xo[s * kRtTile] = static_cast<uint32_t>(m[0] >> 32);
            xo[(k + s) * kRtTile] = static_cast<uint32_t>(c0 + bi);
          }
        }
      }
      xm[0] = mloc;
      xm[kRtTile] = sloc;
      epi_bar_sync<kEpi>();
      // global softmax normaliser from the G partial (max, sum) pairs, in group order
      const float* x0 = xch + tok;
      float mx = x0[0];
#pragma unroll
      for (int g = 1; g < G; ++g) mx = fmaxf(mx, x0[g * gstride]);
      float ssum = 0.f;
#pragma unroll
      for (int g = 0; g < G; ++g) ssum += x0[g * gstride + kRtTile] * expf(x0[g * gstride] - mx);
      const float inv = 1.0f / ssum;
      const float scale = expf(mref - mx) * inv;
      // tile statistics: sum_t p[t, e] per expert (transposed shuffle tree, the four warps in a fixed order later)
      float* st = stat_s + (buf * 4 + quarter) * (EP + 1);
      {
#pragma unroll
        for (int c = 0; c < CPT; ++c) ex[c] = live ? ex[c] * scale : 0.f;
        const float colsum = warp_colsum<CPT>(ex, lane);
        if ((lane & (32 / CPT - 1)) == 0) st[c0 + col_of_lane<CPT>(lane)] = colsum;
      }
      if (grp == 0) {
        const float lse = mx + logf(ssum);
        const float z = warp_sum(live ? lse * lse : 0.f);
        if (lane == 0) st[EP] = z;
        // merge the G sorted candidate lists (ties between groups go to the lower group = lower expert index)
        const uint32_t* q0 = reinterpret_cast<const uint32_t*>(x0) + 2 * kRtTile;
        uint32_t ptrs = 0;
        int sel[8];
        float gate[8];
        float gsum = 0.f;
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          if (s < k) {
            uint32_t best_o = 0;
            int best_g = 0;
#pragma unroll
            for (int g = 0; g < G; ++g) {
              const int pg = (ptrs >> (4 * g)) & 15;
              const uint32_t o = pg < k ? q0[g * gstride + pg * kRtTile] : 0u;
              if (o > best_o) { best_o = o; best_g = g; }
            }
            const int pb = (ptrs >> (4 * best_g)) & 15;
            sel[s] = static_cast<int>(q0[best_g * gstride + (k + pb) * kRtTile]);
            ptrs += 1u << (4 * best_g);
            const float best = __uint_as_float((best_o & 0x80000000u) ? (best_o & 0x7fffffffu) : ~best_o);
            gate[s] = expf(best - mx) * inv;
            gsum += gate[s];
          }
        }
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          if (s < k) {
            // assignment counts: one shared-memory atomic per distinct expert of the warp (match_any), not per token
            const unsigned same = __match_any_sync(0xffffffffu, live ? sel[s] : 64 + lane);
            if (live) {
              a.topk_idx[static_cast<size_t>(t) * k + s] = sel[s];
              a.topk_gate[static_cast<size_t>(t) * k + s] = a.normalize ? gate[s] / (gsum + 1e-9f) : gate[s];
              if (lane == __ffs(same) - 1) atomicAdd(&cnt_s[buf * 512 + s * E + sel[s]], __popc(same));
            }
          }
        }
      }
      epi_bar_sync<kEpi>();
      for (int e = et; e <= E; e += kEpi) {
        const int col = (e == E) ? EP : e;
        const float* sb = stat_s + buf * 4 * (EP + 1);
        const float s = ((sb[col] + sb[(EP + 1) + col]) + sb[2 * (EP + 1) + col]) + sb[3 * (EP + 1) + col];
        a.block_stats[static_cast<size_t>(tile) * (E + 1) + e] = s;
      }
      for (int i = et; i < kE; i += kEpi) {
        a.block_counts[static_cast<size_t>(tile) * kE + i] = cnt_s[buf * 512 + i];
        cnt_s[buf * 512 + i] = 0;                 // this buffer is next used two tiles (two barriers) from now
      }
      buf ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

// Wg fp32 [E,d] -> bf16 [2*EP, d]: rows 0..E-1 = hi, rows EP..EP+E-1 = lo = bf16(w - hi); padding rows zero.
__global__ void router_split_wg_kernel(const float* __restrict__ wg, __nv_bfloat16* __restrict__ out, int E, int EP,
                                       int d) {
  const int n = 2 * EP * d;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int row = i / d, col = i - row * d;
    const int e = row < EP ? row : row - EP;
    float v = 0.f;
    if (e < E) {
      const float w = wg[static_cast<size_t>(e) * d + col];
      const __nv_bfloat16 hi = __float2bfloat16_rn(w);
      v = row < EP ? w : w - __bfloat162float(hi);
    }
    out[i] = __float2bfloat16_rn(v);
  }
}

int make_tmap_bf16_2d(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint32_t box_inner,
                      uint32_t box_outer);   // gemm.cu

template <int EP, int G>
static int launch_router(const void* x, const void* wsplit, RouterArgs a, cudaStream_t st) {
  using Cfg = RouterCfg<EP, G>;
  CUtensorMap tmX, tmW;
  if (int rc = make_tmap_bf16_2d(&tmX, x, a.d, a.T, 64, kRtTile)) return rc;
  if (int rc = make_tmap_bf16_2d(&tmW, wsplit, a.d, 2 * EP, 64, 2 * EP)) return rc;
  auto kern = router_fwd_kernel<EP, G>;
  static DeviceOnce once;
  const int dev = current_device();
  if (!once.done(dev)) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kRtSmemBudget / Cfg::kCtasPerSm);
    B200MOE_REQUIRE(e == cudaSuccess, B200MOE_ECUDA, "router_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    once.set(dev);
  }
  a.stages = Cfg::stages(a.k);
  static const int stage_cap = [] { const char* e = getenv("B200MOE_ROUTER_STAGES"); return e ? atoi(e) : 0; }();
  static const int cta_cap = [] { const char* e = getenv("B200MOE_ROUTER_CTAS"); return e ? atoi(e) : 0; }();
  if (stage_cap >= 2 && stage_cap < a.stages) a.stages = stage_cap;
  B200MOE_REQUIRE(a.stages >= 2, B200MOE_EINVAL, "router_fwd: no shared memory left for the load ring (top_k=%d)", a.k);
  const int ntiles = (a.T + kRtTile - 1) / kRtTile;
  const int max_ctas = cta_cap > 0 ? cta_cap : num_sms() * Cfg::kCtasPerSm;
  const int grid = ntiles < max_ctas ? ntiles : max_ctas;
  launch_k(kern, grid, Cfg::kThreads, Cfg::total(a.k), st, tmX, tmW, a);
  return check_launch("router_fwd");
}

static int router_ep(int E) { return E <= 16 ? 16 : (E <= 32 ? 32 : 64); }

}  // namespace b200moe

extern "C" int64_t b200moe_router_scratch_bytes(int d, int E) {
  if (d <= 0 || E < 1 || E > 64) return -1;
  return static_cast<int64_t>(2) * b200moe::router_ep(E) * d * 2;
}

static int router_fwd_impl(const void* x, const float* ctx_bias, const b200moe_ctx_desc* ctx, int tokens_per_image,
                           const float* wg, void* wg_split, int T, int d, int E, int k, int normalize_topk,
                           float* logits, int32_t* topk_idx, float* topk_gate, int32_t* block_counts,
                           float* block_stats, void* stream) {
  using namespace b200moe;
  B200MOE_REQUIRE(T >= 0 && d > 0 && d % 64 == 0, B200MOE_EINVAL, "router_fwd: d=%d must be a positive multiple of 64", d);
  B200MOE_REQUIRE(E >= 1 && E <= 64, B200MOE_EINVAL, "router_fwd: num_experts=%d outside [1,64]", E);
  B200MOE_REQUIRE(k >= 1 && k <= E && k <= 8, B200MOE_EINVAL, "router_fwd: top_k=%d must be in [1,min(E,8)]", k);
  B200MOE_REQUIRE((ctx_bias == nullptr && ctx == nullptr) || tokens_per_image > 0, B200MOE_EINVAL,
                  "router_fwd: tokens_per_image <= 0");
  B200MOE_REQUIRE(wg_split != nullptr, B200MOE_EINVAL, "router_fwd: wg_split scratch is NULL (b200moe_router_scratch_bytes)");
  CtxArgs ca{};
  if (ctx != nullptr) {
    B200MOE_REQUIRE(ctx_bias == nullptr, B200MOE_EINVAL, "router_fwd: pass ctx_bias or ctx, not both");
    B200MOE_REQUIRE(tokens_per_image >= kRtMinTokPerImg, B200MOE_EINVAL,
                    "router_fwd: in-kernel context bias needs tokens_per_image >= %d (got %d): use b200moe_ctx_bias_fwd",
                    kRtMinTokPerImg, tokens_per_image);
    B200MOE_REQUIRE(ctx->context && ctx->wc && ctx->n_fields >= 1 && ctx->n_fields <= B200MOE_MAX_CTX_FIELDS &&
                        ctx->context_dim > 0 && ctx->n_fields * ctx->context_dim <= kRtMaxNcd, B200MOE_EINVAL,
                    "router_fwd: bad context descriptor (n_fields * context_dim must be <= %d for the in-kernel bias)", kRtMaxNcd);
    ca.context = ctx->context; ca.wc = ctx->wc; ca.n = ctx->n_fields; ca.cd = ctx->context_dim;
    for (int i = 0; i < B200MOE_MAX_CTX_FIELDS; ++i) {
      ca.tb.table[i] = i < ctx->n_fields ? ctx->tables[i] : nullptr;
      ca.tb.dtable[i] = nullptr;
      ca.tb.vocab[i] = i < ctx->n_fields ? ctx->vocab[i] : 0;
      B200MOE_REQUIRE(i >= ctx->n_fields || (ctx->tables[i] != nullptr && ctx->vocab[i] > 0), B200MOE_EINVAL,
                      "router_fwd: bad context table %d", i);
    }
  }
  if (T == 0) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (tokens_per_image <= 0) tokens_per_image = T;
  const int EP = router_ep(E);
  if (wg != nullptr) {   // NULL: wg_split was already filled by b200moe_prepare_weights
    const int n = 2 * EP * d;
    int blocks = (n + 255) / 256;
    if (blocks > 4 * num_sms()) blocks = 4 * num_sms();
    router_split_wg_kernel<<<blocks, 256, 0, st>>>(wg, static_cast<__nv_bfloat16*>(wg_split), E, EP, d);
    if (int rc = check_launch("router_split_wg")) return rc;
  }
  RouterArgs a{ca, ctx_bias, tokens_per_image, T, d, E, k, normalize_topk, logits, topk_idx, topk_gate, block_counts,
               block_stats, 0};
  if (EP == 16) return E <= 8 ? launch_router<16, 1>(x, wg_split, a, st) : launch_router<16, 2>(x, wg_split, a, st);
  if (EP == 32) return launch_router<32, 2>(x, wg_split, a, st);
  return launch_router<64, 4>(x, wg_split, a, st);
}

extern "C" int b200moe_router_fwd(const void* x, const float* ctx_bias, int tokens_per_image, const float* wg,
                                  void* wg_split, int T, int d, int E, int k, int normalize_topk, float* logits,
                                  int32_t* topk_idx, float* topk_gate, int32_t* block_counts, float* block_stats,
                                  void* stream) {
  return router_fwd_impl(x, ctx_bias, nullptr, tokens_per_image, wg, wg_split, T, d, E, k, normalize_topk, logits,
                         topk_idx, topk_gate, block_counts, block_stats, stream);
}

extern "C" int b200moe_router_fwd_ctx(const void* x, const b200moe_ctx_desc* ctx, int tokens_per_image, void* wg_split,
                                      int T, int d, int E, int k, int normalize_topk, float* logits, int32_t* topk_idx,
                                      float* topk_gate, int32_t* block_counts, float* block_stats, void* stream) {
  return router_fwd_impl(x, nullptr, ctx, tokens_per_image, nullptr, wg_split, T, d, E, k, normalize_topk, logits,
                         topk_idx, topk_gate, block_counts, block_stats, stream);
}
