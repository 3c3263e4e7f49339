// Grouped expert GEMMs on tcgen05 / TMEM fed by TMA (sm_100a).  SURVEY.md Appendix A.6 and §8(d).
//
// One persistent warp-specialised kernel template covers the six contractions of the expert FFN:
//   row mode  (M = rows of the permuted buffer, 128-row tiles never straddle experts; tile_expert[] picks the
//              weight slab): up-proj, down-proj (A and B K-major), dgrad-down, dgrad-up (B = weights read MN-major,
//              i.e. W[e] is used as [K,N] without a transposed copy);
//   wgrad mode (M,N = weight dims, K = rows of one expert): A and B are both activations read MN-major.
// Roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (one elected lane),
// warps 2..9 = epilogue (warp w owns TMEM lanes 32*(w%4)..+31 = output rows and one half of the tile's columns;
// results leave through swizzled smem boxes and TMA stores, fp32 weight gradients optionally as TMA reduce-add).
// A kStages-deep smem ring
// (full/empty mbarriers) decouples TMA from MMA; two TMEM accumulator stages (tmem_full/tmem_empty) let the
// epilogue of tile i overlap the main loop of tile i+1.
//
// Shared-memory operand tiles are 128-byte-swizzled exactly as TMA writes them:
//   K-major  tile [rows][64 bf16]      : 8-row groups 1024 B apart (SBO); one k-step (16 bf16) = +32 B
//   MN-major tile [64 k-rows][64 bf16] per 64-wide chunk of M/N, chunks 8192 B apart (LBO), 8-k-row groups
//            1024 B apart (SBO); one k-step (16 k-rows) = +2048 B
#include "common.cuh"
#include "postnorm.cuh"

namespace b200moe {

// EPI_DOWN_COMBINE: the down-projection with the gate-weighted combine of A.7 fused into its epilogue (y accumulated in
// fp32 straight from the accumulator registers; the expert outputs yo are stored only when the backward needs them).
enum : int { EPI_UP = 0, EPI_DOWN = 1, EPI_DGRAD_DOWN = 2, EPI_DGRAD_UP = 3, EPI_WGRAD = 4, EPI_DOWN_COMBINE = 5 };
enum : int { ACT_GELU = 0, ACT_RELU = 1 };   // encoder FFN (hybrid_encoder.py:157, utils.py:160-161) / decoder FFN (rtdetrv2_decoder.py:169)

constexpr int BM = 128;
constexpr int BK = 64;
// Epilogue warps: the bf16 epilogues (bias / GELU / GELU' / pack / TMA store) are latency chains (TMEM load, proxy
// fence, bulk-store read-back), so they run 16 warps = four per scheduler, four per TMEM lane quarter, each taking a
// quarter of the tile's columns through a private 2 KB staging box.  The fp32 weight-gradient epilogue is a plain
// copy: 8 warps with 4 KB boxes.
__host__ __device__ constexpr int epi_warps(int epi) { return epi == 4 /*EPI_WGRAD*/ ? 8 : 16; }
__host__ __device__ constexpr int gemm_threads(int epi) { return 64 + 32 * epi_warps(epi); }
// (Tried in round 2 and dropped: two staging boxes per warp for the fp32 weight-gradient copy-out, paid for with one ring
// stage -- 64 experts x 512 rows, d=1024, F=4096: 598 -> 633 us; the kernel already writes its 2.1 GB at 3.6 TB/s next to
// 0.7 GB of reads, i.e. it sits on the HBM read+write limit, and the shallower ring costs more than the second box gains.)
// (Tried in round 2 and dropped: a two-stage, 8-epilogue-warp configuration with TWO CTAs per SM for the latency-bound
// problem sizes.  At 3200 tokens every GEMM got slower -- up 11.9 -> 14.4 us, dgrad_down 12.1 -> 18.4 us: two k-blocks in
// flight do not cover the TMA latency -- and the co-residency of independent branches did not make up for it: 104 -> 107
// us per step, gpurun_out timelines of round 2.)

struct GemmArgs {
  const int32_t* meta;            // row mode: meta[1] = number of 128-row tiles in use
  const int32_t* tile_expert;     // row mode
  const int32_t* expert_offsets;  // wgrad mode
  int K;        // row mode: reduction length (d or F)
  int N;        // output columns per expert
  int M;        // wgrad mode: output rows per expert
  int E;
  const float* bias;              // [E,N] fp32 or null
  const __nv_bfloat16* aux;       // EPI_DGRAD_DOWN: pre [R,N]
  __nv_bfloat16* out0;            // bf16 output [R,N]
  __nv_bfloat16* out1;            // EPI_UP: second output (pre) or null
  float* outf;                    // EPI_WGRAD: [E,M,N]
  int accumulate;
  int act;                        // EPI_UP / EPI_DGRAD_DOWN: ACT_GELU or ACT_RELU
  // EPI_DOWN_COMBINE
  const int32_t* token_of_row;       // [R], -1 = pad row
  const float* gate_of_row;          // [R]
  float* yacc;                       // [T,N] fp32, zero before the launch (reduction mode), or null (gather mode)
  const int32_t* row_of_assignment;  // [T,k]  gather mode: the tail combines from the stored yo rows
  const float* topk_gate;            // [T,k]
  int topk;
  int32_t* grid_sync;                // one counter, zero before the launch, or null: the caller finishes y from yacc
  void* yout;                        // [T,N]: written from yacc by the kernel's tail once every CTA has delivered --
  int T;                             //        bf16 y, or with ln_gamma the fused post-norm LayerNorm(residual + y)
  const void* ln_res;                // [T,N] residual (bf16 / fp32), with ln_gamma
  const float* ln_gamma;             // [N] or null: plain conversion
  const float* ln_beta;
  float* ln_mean_rstd;               // [T,2]
  float ln_eps;
  int ln_res_f32, ln_out_f32;
  // A-stationary schedule (template ARES): the A tile of a unit (= `nrun` consecutive N tiles of one M tile) is loaded
  // once and stays in shared memory, only B streams through the ring
  int nrun;                          // N tiles per unit (divides N / BN); ignored without ARES
  int nstages;                       // ring depth with ARES (<= K / BK and <= 8)
  // operand rows fetched through a row map by TMA (tile::gather4) instead of read from a permuted copy:
  // row mode: row r of A is row gather_idx[r] of the tensor behind tmA; wgrad mode: the same for the k-rows of B.
  // Negative indices deliver zeros (pad rows).
  const int32_t* gather_idx;
  // wgrad mode, second problem of the same launch (both weight gradients of the FFN: their tiles share one persistent
  // grid, so neither ends on a partly filled wave and the second does not queue behind the first): operands behind
  // tmA2 / tmB2, output [E,M2,N2] behind tmC1; M2 == 0: one problem
  int M2, N2;
  int32_t* zero_i32;                 // one int the kernel clears (the rendezvous counter of the fused combine that follows)
};

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <int BN, bool CTA2>
struct SmemCfg {
  static constexpr int kBLoad = CTA2 ? BN / 2 : BN;                        // B rows / columns this CTA loads
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = kBLoad * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // (tried in round 2: 4 stages for the BN = 128 tiles, so that their 160 KB footprint leaves room for the kernels of the
  // other branches beside the persistent CTA -- isolated GEMMs at 3200 tokens 16.7 -> 18.9 us, step time unchanged)
  static constexpr int kStages = (kStageBytes > 32768) ? 4 : 6;
  static constexpr int kOutBytes = 8 * 4096;                               // epilogue staging: 8 x 4 KB or 16 x 2 KB
  static constexpr int kBarBytes = 512;
  static constexpr int kTotal = kStages * kStageBytes + kOutBytes + kBarBytes + 1024;  // + alignment slack
};

// CTA2: clusters of two CTAs share one 256 x BN tile (tcgen05 cta_group::2).  Each CTA loads its own 128 rows of A and
// HALF of the B tile, CTA 0 issues one M=256 MMA that reads both halves, and each CTA's TMEM receives its own 128
// output rows: per FLOP the pair pulls 2/3 of the operand bytes through L2 that two independent CTAs would -- the
// 128 x 256 tiles are L2-bandwidth-bound (ncu: ~14 TB/s of L2->SM traffic at 58-72 % tensor-pipe activity).
// ARES (row mode, K = d_model <= 512): A-stationary.  The 128 x 256 tiles of the up-projection and of dgrad-down stream
// 48 KB per k-block through L2 for 512 tensor-pipe cycles -- 14 KB per clock over the chip against the ~6.3 KB per clock
// the L2 slices deliver (ncu round 2: 11-12 TB/s of TMA loads at 44 % tensor-pipe activity).  With the A tile resident
// ([K/64] boxes of 16 KB) and CTA pairs splitting B, a k-block costs 16 KB per CTA: a third of the traffic.
template <int BN, bool A_MN, bool B_MN, int EPI, bool CTA2, bool ARES = false>
__global__ void __launch_bounds__(gemm_threads(EPI), 1)
grouped_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC0, const __grid_constant__ CUtensorMap tmC1,
                    const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                    const GemmArgs args) {
  using Cfg = SmemCfg<BN, CTA2>;
  constexpr int kStages = Cfg::kStages;
  constexpr int kMaxStages = ARES ? 8 : kStages;
  constexpr int kRingStage = ARES ? Cfg::kBBytes : Cfg::kStageBytes;      // bytes of one ring slot
  constexpr bool kWgrad = (EPI == EPI_WGRAD);
  static_assert(!ARES || (!kWgrad && !A_MN), "the A-stationary schedule is a row-mode schedule");
  constexpr int kBLoad = Cfg::kBLoad;
  constexpr int kTileM = CTA2 ? 2 * BM : BM;                               // rows of one scheduled tile
  constexpr int kEpiWarps = epi_warps(EPI);
  constexpr int kEpiThreads = kEpiWarps * 32;
  constexpr int kStgBytes = Cfg::kOutBytes / kEpiWarps;                    // staging per warp: 4 KB, or 2 KB with 16 warps
  constexpr int kBoxBytes = kWgrad ? 4096 : 2048;                          // one [32 x 32] box: fp32 or bf16
  constexpr uint32_t kTmemCols = 2 * BN;  // 256 or 512: power of two
  constexpr uint32_t kIdesc = make_idesc_bf16(kTileM, BN, A_MN, B_MN);

  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int nstages = ARES ? args.nstages : kStages;
  unsigned char* a_res = smem;                                             // ARES: [K / BK] resident A boxes
  unsigned char* ring = ARES ? smem + (args.K / BK) * Cfg::kABytes : smem;
  unsigned char* stage_out = ring + nstages * kRingStage;                  // per-warp epilogue staging boxes
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(stage_out + Cfg::kOutBytes);
  uint64_t* empty_bar = full_bar + kMaxStages;
  uint64_t* tfull_bar = empty_bar + kMaxStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  uint64_t* ld_bar = reinterpret_cast<uint64_t*>(tmem_slot + 2);           // EPI_DGRAD_DOWN: one per epilogue warp

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = CTA2 ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC0);
    if (EPI == EPI_UP || EPI == EPI_DGRAD_DOWN) tma_prefetch_desc(&tmC1);
    if (kWgrad && args.M2 != 0) { tma_prefetch_desc(&tmA2); tma_prefetch_desc(&tmB2); tma_prefetch_desc(&tmC1); }
    if (EPI == EPI_DGRAD_DOWN)
      for (int w = 0; w < kEpiWarps; ++w) mbar_init(&ld_bar[w], 1);
    for (int s = 0; s < nstages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], (CTA2 ? 2 : 1) * kEpiThreads); }
    fence_mbar_init();
  }
  if (warp == 1) {
    if (CTA2) tmem_alloc2<kTmemCols>(tmem_slot);
    else tmem_alloc<kTmemCols>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  if (CTA2) cluster_sync_all();                 // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_sync();                                   // prologue done; everything below reads the previous kernel's output

  // ---- tile schedule (identical in every role) -------------------------------------------------------------
  const int tiles_n = args.N / BN;
  const int tiles_m = kWgrad ? args.M / kTileM : 0;
  const int tiles_n2 = kWgrad ? args.N2 / BN : 0;                          // second weight gradient of the launch (or 0)
  const int tiles_m2 = kWgrad ? args.M2 / kTileM : 0;
  const int total_tiles = kWgrad ? args.E * (tiles_m * tiles_n + tiles_m2 * tiles_n2)
                                 : (args.meta[1] / (kTileM / BM)) * tiles_n;
  const int worker = CTA2 ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int nworkers = CTA2 ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  // a worker takes whole units: `nrun` consecutive tiles (one M tile, a run of N tiles); nrun == 1 without ARES
  const int nrun = ARES ? args.nrun : 1;
  const int total_units = total_tiles / nrun;
#define FOR_EACH_TILE(tile)                                              \
  for (int unit_ = worker; unit_ < total_units; unit_ += nworkers)        \
    for (int j_ = 0, tile = unit_ * nrun; j_ < nrun; ++j_, ++tile)
  if (args.zero_i32 != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *args.zero_i32 = 0;

  int prob = 0;                                     // wgrad mode: which of the launch's two problems the tile belongs to
  auto decode = [&](int tile, int& e, int& m_tile, int& n_tile, int& krow0, int& kblocks) {
    if (kWgrad) {
      prob = (tiles_m2 != 0 && tile >= args.E * tiles_m * tiles_n) ? 1 : 0;
      if (prob) tile -= args.E * tiles_m * tiles_n;
      const int tn = prob ? tiles_n2 : tiles_n;
      const int per_e = (prob ? tiles_m2 : tiles_m) * tn;
      e = tile / per_e;
      const int r = tile - e * per_e;
      m_tile = r / tn;
      n_tile = r - m_tile * tn;
      if (CTA2) m_tile = 2 * m_tile + static_cast<int>(cta_rank);         // this CTA's 128 rows of the 256-row tile
      const int r0 = args.expert_offsets[e], r1 = args.expert_offsets[e + 1];
      krow0 = r0;
      kblocks = (r1 - r0) / BK;
    } else {
      m_tile = tile / tiles_n;
      n_tile = tile - m_tile * tiles_n;
      if (CTA2) m_tile = 2 * m_tile + static_cast<int>(cta_rank);         // both 128-row tiles belong to one expert
      e = args.tile_expert[m_tile];
      krow0 = 0;
      kblocks = args.K / BK;
    }
  };

  if (warp == 0) {
    // ======================= TMA producer =======================
    // One elected lane issues the tiled loads.  With a row map (args.gather_idx) the whole warp takes part: a
    // tile::gather4 load moves four rows, lane l fetches rows 4l .. 4l+3 of a 128-row K-major box, or (wgrad, B read
    // MN-major) lanes 0-15 / 16-31 fetch the 64 k-rows of alternate 64-column chunks.
    const bool gather = args.gather_idx != nullptr;
    if (lane == 0 || gather) {
      int stage = 0;
      uint32_t phase = 0;
      auto a_row0 = [&](int unit) {                       // first A row this CTA loads for a unit (row mode)
        const int m = (unit * nrun) / tiles_n;
        return (CTA2 ? 2 * m + static_cast<int>(cta_rank) : m) * BM;
      };
      int4 gi_next = make_int4(-1, -1, -1, -1);           // row map of the next unit, fetched one unit ahead
      if (gather && !kWgrad && worker < total_units)
        gi_next = *reinterpret_cast<const int4*>(args.gather_idx + a_row0(worker) + 4 * lane);
      for (int unit_ = worker; unit_ < total_units; unit_ += nworkers) {
        const int4 gi = gi_next;
        if (gather && !kWgrad && unit_ + nworkers < total_units)
          gi_next = *reinterpret_cast<const int4*>(args.gather_idx + a_row0(unit_ + nworkers) + 4 * lane);
        for (int j_ = 0, tile = unit_ * nrun; j_ < nrun; ++j_, ++tile) {
          int e, m_tile, n_tile, krow0, kblocks;
          decode(tile, e, m_tile, n_tile, krow0, kblocks);
          const bool load_a = !ARES || j_ == 0;           // A-stationary: the unit's first tile brings A in
          int4 bi = make_int4(-1, -1, -1, -1);            // wgrad: row map of the k-block, fetched one k-block ahead
          if (gather && kWgrad && kblocks > 0)
            bi = *reinterpret_cast<const int4*>(args.gather_idx + krow0 + 4 * (lane & 15));
          for (int kb = 0; kb < kblocks; ++kb) {
            const int4 bi_cur = bi;
            if (gather && kWgrad && kb + 1 < kblocks)
              bi = *reinterpret_cast<const int4*>(args.gather_idx + krow0 + (kb + 1) * BK + 4 * (lane & 15));
            mbar_wait(&empty_bar[stage], phase ^ 1);
            unsigned char* sa = ARES ? a_res + kb * Cfg::kABytes : ring + stage * kRingStage;
            unsigned char* sb = ARES ? ring + stage * kRingStage : sa + Cfg::kABytes;
            // the bytes of BOTH CTAs land on CTA 0's barrier, armed by CTA 0 alone
            if (lane == 0 && leader)
              mbar_arrive_expect_tx(&full_bar[stage], (CTA2 ? 2 : 1) * ((load_a ? Cfg::kABytes : 0) + Cfg::kBBytes));
            if (gather) __syncwarp();
            auto load = [&](void* dst, const CUtensorMap* tm, int c0, int c1) {
              if (CTA2) tma_load_2d_pair(dst, tm, &full_bar[stage], c0, c1);
              else tma_load_2d(dst, tm, &full_bar[stage], c0, c1);
            };
            auto load4 = [&](void* dst, const CUtensorMap* tm, int c0, const int4& rows) {
              if (CTA2) tma_gather4_pair(dst, tm, &full_bar[stage], c0, rows);
              else tma_gather4(dst, tm, &full_bar[stage], c0, rows);
            };
            if (load_a) {
              if (A_MN) {
                if (lane == 0) {
                  const CUtensorMap* pA = (kWgrad && prob) ? &tmA2 : &tmA;
#pragma unroll
                  for (int j = 0; j < BM / 64; ++j) load(sa + j * 8192, pA, m_tile * BM + j * 64, krow0 + kb * BK);
                }
              } else if (gather) {
                load4(sa + lane * 512, &tmA, kb * BK, gi);          // rows 4l .. 4l+3 of the box, 128 B each
              } else if (lane == 0) {
                load(sa, &tmA, kb * BK, m_tile * BM);
              }
            }
            const int n0 = n_tile * BN + static_cast<int>(cta_rank) * kBLoad;      // this CTA's share of the B tile
            if (B_MN) {
              const int krow = kWgrad ? krow0 + kb * BK : e * args.K + kb * BK;
              if (kWgrad && gather) {
                for (int j = lane >> 4; j < kBLoad / 64; j += 2)
                  load4(sb + j * 8192 + (lane & 15) * 512, &tmB, n0 + j * 64, bi_cur);
              } else if (lane == 0) {
                const CUtensorMap* pB = (kWgrad && prob) ? &tmB2 : &tmB;
#pragma unroll
                for (int j = 0; j < kBLoad / 64; ++j) load(sb + j * 8192, pB, n0 + j * 64, krow);
              }
            } else if (lane == 0) {
              load(sb, &tmB, kb * BK, e * args.N + n0);
            }
            if (++stage == nstages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================= MMA issuer =======================
    if (lane == 0 && leader) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      FOR_EACH_TILE(tile) {
        int e, m_tile, n_tile, krow0, kblocks;
        decode(tile, e, m_tile, n_tile, krow0, kblocks);
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(ARES ? a_res + kb * Cfg::kABytes : ring + stage * kRingStage);
          const uint32_t sb = ARES ? smem_u32(ring + stage * kRingStage) : sa + Cfg::kABytes;
          const uint64_t da = A_MN ? make_smem_desc(sa, 8192, 1024) : make_smem_desc(sa, 16, 1024);
          const uint64_t db = B_MN ? make_smem_desc(sb, 8192, 1024) : make_smem_desc(sb, 16, 1024);
#pragma unroll
          for (int ks = 0; ks < BK / 16; ++ks) {
            const uint64_t a_k = da + static_cast<uint64_t>(ks * (A_MN ? 128 : 2));
            const uint64_t b_k = db + static_cast<uint64_t>(ks * (B_MN ? 128 : 2));
            if (CTA2) umma_bf16_ss_pair(tmem_d, a_k, b_k, kIdesc, (kb | ks) != 0 ? 1u : 0u);
            else umma_bf16_ss(tmem_d, a_k, b_k, kIdesc, (kb | ks) != 0 ? 1u : 0u);
          }
          // frees the smem slot (of both CTAs) once these MMAs have read it
          if (CTA2) umma_commit_pair(&empty_bar[stage]);
          else umma_commit(&empty_bar[stage]);
          if (++stage == nstages) { stage = 0; phase ^= 1; }
        }
        if (kblocks > 0) {                               // accumulator complete (in both CTAs' TMEM)
          if (CTA2) umma_commit_pair(&tfull_bar[acc]);
          else umma_commit(&tfull_bar[acc]);
        } else {                                         // empty expert: nothing to wait for
          mbar_arrive(&tfull_bar[acc]);
          if (CTA2) mbar_arrive_remote(&tfull_bar[acc], 1);
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ======================= epilogue (warps 2 ..) =======================
    // TMEM -> registers -> (bias / GELU / GELU') -> swizzled smem box -> TMA store.  Each warp owns 32 output rows
    // and a private staging box, so no cross-warp synchronisation is needed: lane 0 issues the bulk stores
    // and is the one that waits for them to have read the staging box before it is overwritten.
    const int quarter = warp & 3;              // TMEM lane quarter this warp may read
    const int cgroup = (warp - 2) >> 2;        // which share of the tile's columns
    constexpr int kChunksPerWarp = (BN / 32) / (kEpiWarps / 4);
    constexpr bool kCombine = (EPI == EPI_DOWN_COMBINE);
    unsigned char* stg = stage_out + (warp - 2) * kStgBytes;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t ld_phase = 0;
    FOR_EACH_TILE(tile) {
      int e, m_tile, n_tile, krow0, kblocks;
      decode(tile, e, m_tile, n_tile, krow0, kblocks);
      const int row0 = m_tile * BM + quarter * 32;                 // first of this warp's 32 output rows
      const size_t grow = static_cast<size_t>(row0) + lane;        // this thread's output row
      // EPI_DGRAD_DOWN: the pre-activation box of a chunk travels by TMA into this warp's staging box (the same box the
      // result leaves from; a thread reads and writes only its own row of it).  Per-thread 16-byte loads of 64-byte row
      // pieces cost 8 load instructions x 32 lines per chunk and kept the SM's load/store unit busy; the first box is
      // requested before the wait for the accumulator.
      if (EPI == EPI_DGRAD_DOWN && lane == 0) {
        tma_store_wait_read0();
        mbar_arrive_expect_tx(&ld_bar[warp - 2], kBoxBytes);
        tma_load_2d(stg, &tmC1, &ld_bar[warp - 2], n_tile * BN + cgroup * kChunksPerWarp * 32, row0);
      }
      // bias of this warp's chunks, one column per lane (a 128-byte line per chunk), requested before the wait for the
      // accumulator and broadcast by shuffles below: the per-chunk broadcast loads (8 x LDG.128 behind the TMEM load)
      // put an L2 round trip into every chunk's latency chain
      float bias_lane[kChunksPerWarp];
      if (EPI == EPI_UP || EPI == EPI_DOWN || kCombine) {
#pragma unroll
        for (int cc = 0; cc < kChunksPerWarp; ++cc)
          bias_lane[cc] = __ldg(args.bias + static_cast<size_t>(e) * args.N + n_tile * BN +
                                (cgroup * kChunksPerWarp + cc) * 32 + lane);
      }
      // fused combine: this thread's row belongs to token `tok` (-1: pad row) and carries the gate `gate`
      int tok = -1;
      float gate = 0.f;
      if (kCombine && args.yacc != nullptr) {
        tok = args.token_of_row[grow];
        gate = args.gate_of_row[grow];
      }
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int cc = 0; cc < kChunksPerWarp; ++cc) {
        const int c = cgroup * kChunksPerWarp + cc;
        const int col0 = n_tile * BN + c * 32;
        uint32_t v[32];
        tmem_ld32(taddr + c * 32, v);
        uint4 pr[4];
        if (EPI == EPI_DGRAD_DOWN) {
          mbar_wait(&ld_bar[warp - 2], ld_phase);
          ld_phase ^= 1;
          const uint32_t swz_l = (lane >> 1) & 3;
#pragma unroll
          for (int j = 0; j < 4; ++j) pr[j] = *reinterpret_cast<const uint4*>(stg + lane * 64 + ((j ^ swz_l) << 4));
        }
        tmem_ld_wait();
        if (EPI == EPI_WGRAD) {
          // fp32 box [32 rows][32 cols] = 4 KB, 128-byte swizzle: 16-byte chunk j of row r sits at j ^ (r & 7)
          if (lane == 0) tma_store_wait_read0();
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint4 q = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            if (kblocks == 0) q = make_uint4(0, 0, 0, 0);
            *reinterpret_cast<uint4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) = q;
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            const int grow0 = e * (prob ? args.M2 : args.M) + row0;
            const CUtensorMap* pC = prob ? &tmC1 : &tmC0;
            if (args.accumulate) tma_reduce_add_2d(pC, stg, col0, grow0);
            else tma_store_2d(pC, stg, col0, grow0);
            tma_store_commit();
          }
        } else {
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          if (EPI == EPI_UP || EPI == EPI_DOWN || kCombine) {
            float bl = bias_lane[0];
#pragma unroll
            for (int q = 1; q < kChunksPerWarp; ++q) bl = (cc == q) ? bias_lane[q] : bl;
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] += __shfl_sync(0xffffffffu, bl, j);
          }
          // bf16 box [32 rows][32 cols] = 2 KB, 64-byte swizzle: 16-byte chunk j of row r sits at j ^ ((r >> 1) & 3)
          const uint32_t swz = (lane >> 1) & 3;
          if (EPI == EPI_UP && args.out1) {    // pre-activation leaves first; its bulk store reads the box during the GELU math
            if (lane == 0) tma_store_wait_read0();
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 pk;
              pk.x = float2_to_bf16x2(f[8 * j], f[8 * j + 1]); pk.y = float2_to_bf16x2(f[8 * j + 2], f[8 * j + 3]);
              pk.z = float2_to_bf16x2(f[8 * j + 4], f[8 * j + 5]); pk.w = float2_to_bf16x2(f[8 * j + 6], f[8 * j + 7]);
              *reinterpret_cast<uint4*>(stg + lane * 64 + ((j ^ swz) << 4)) = pk;
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmC1, stg, col0, row0);
              tma_store_commit();
            }
          }
          if (EPI == EPI_UP) {
            if (args.act == ACT_RELU) {
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = gelu_erf(f[j]);
            }
          }
          if (EPI == EPI_DGRAD_DOWN) {
            if (args.act == ACT_RELU) {      // d relu(v)/dv = [v > 0] (0 at v == 0, as torch)
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                float2 t;
                t = bf16x2_to_float2(pr[j].x); f[8 * j] = t.x > 0.f ? f[8 * j] : 0.f; f[8 * j + 1] = t.y > 0.f ? f[8 * j + 1] : 0.f;
                t = bf16x2_to_float2(pr[j].y); f[8 * j + 2] = t.x > 0.f ? f[8 * j + 2] : 0.f; f[8 * j + 3] = t.y > 0.f ? f[8 * j + 3] : 0.f;
                t = bf16x2_to_float2(pr[j].z); f[8 * j + 4] = t.x > 0.f ? f[8 * j + 4] : 0.f; f[8 * j + 5] = t.y > 0.f ? f[8 * j + 5] : 0.f;
                t = bf16x2_to_float2(pr[j].w); f[8 * j + 6] = t.x > 0.f ? f[8 * j + 6] : 0.f; f[8 * j + 7] = t.y > 0.f ? f[8 * j + 7] : 0.f;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                float2 t;
                t = bf16x2_to_float2(pr[j].x); f[8 * j] *= gelu_erf_grad(t.x); f[8 * j + 1] *= gelu_erf_grad(t.y);
                t = bf16x2_to_float2(pr[j].y); f[8 * j + 2] *= gelu_erf_grad(t.x); f[8 * j + 3] *= gelu_erf_grad(t.y);
                t = bf16x2_to_float2(pr[j].z); f[8 * j + 4] *= gelu_erf_grad(t.x); f[8 * j + 5] *= gelu_erf_grad(t.y);
                t = bf16x2_to_float2(pr[j].w); f[8 * j + 6] *= gelu_erf_grad(t.x); f[8 * j + 7] *= gelu_erf_grad(t.y);
              }
            }
          }
          uint4 pk[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            pk[j].x = float2_to_bf16x2(f[8 * j], f[8 * j + 1]); pk[j].y = float2_to_bf16x2(f[8 * j + 2], f[8 * j + 3]);
            pk[j].z = float2_to_bf16x2(f[8 * j + 4], f[8 * j + 5]); pk[j].w = float2_to_bf16x2(f[8 * j + 6], f[8 * j + 7]);
          }
          if (kCombine && tok >= 0 && args.yacc != nullptr) {
            // y[tok, col0 .. col0+31] += gate * yo[row]: the term is formed from the bf16-rounded expert output (the value
            // the separate combine kernel would read back), added with fp32 reductions at L2.  Two terms commute, so
            // for top_k <= 2 the result does not depend on which tile arrives first.
            float* dst = args.yacc + static_cast<size_t>(tok) * args.N + col0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float v8[8];
              unpack8(pk[j], v8);
              red_add_v4(dst + 8 * j, combine_term(gate, v8[0]), combine_term(gate, v8[1]), combine_term(gate, v8[2]),
                         combine_term(gate, v8[3]));
              red_add_v4(dst + 8 * j + 4, combine_term(gate, v8[4]), combine_term(gate, v8[5]), combine_term(gate, v8[6]),
                         combine_term(gate, v8[7]));
            }
          }
          unsigned char* stg0 = stg;
          if (!kCombine || args.out0 != nullptr) {
          if (lane == 0) tma_store_wait_read0();
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(stg0 + lane * 64 + ((j ^ swz) << 4)) = pk[j];
          fence_proxy_async_smem();
          __syncwarp();
          }
          if (lane == 0 && (!kCombine || args.out0 != nullptr)) {
            tma_store_2d(&tmC0, stg0, col0, row0);
            tma_store_commit();
            if (EPI == EPI_DGRAD_DOWN && cc + 1 < kChunksPerWarp) {   // next pre-activation box, once the store has read this one
              tma_store_wait_read0();
              mbar_arrive_expect_tx(&ld_bar[warp - 2], kBoxBytes);
              tma_load_2d(stg, &tmC1, &ld_bar[warp - 2], col0 + 32, row0);
            }
          }
        }
      }
      tc_fence_before();
      if (CTA2 && !leader) mbar_arrive_remote(&tempty_bar[acc], 0);   // the MMA issuer lives in CTA 0
      else mbar_arrive(&tempty_bar[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (lane == 0) tma_store_wait_all0();      // staging smem must outlive the bulk stores that read it
  }

  tc_fence_before();
  __syncthreads();
  if (CTA2) cluster_sync_all();                 // neither CTA leaves while the other may still touch its smem / TMEM
  if (warp == 1) {                              // TMEM goes back before the rendezvous below: a CTA of this grid that is
    tc_fence_after();                           // still waiting for columns must never wait for CTAs that wait for it
    if (CTA2) tmem_dealloc2<kTmemCols>(tmem_base);
    else tmem_dealloc<kTmemCols>(tmem_base);
  }
  if (EPI == EPI_DOWN_COMBINE && args.grid_sync != nullptr) {
    // Tail of the fused combine: once EVERY CTA of the grid has delivered its rows (one counter; the grid is never larger
    // than what is resident at once, and dependents of a programmatic launch are scheduled only after all CTAs of this
    // grid run), the CTAs share the combine itself -- gather mode: y[t] = sum_s gate * yo[row] from the rows just stored
    // (any top_k, slot order); reduction mode: the fp32 accumulator the epilogues added to is complete in L2 and is
    // converted.  No combine launch follows; tokens without a kept assignment come out as zeros either way.
    if (threadIdx.x == 0) {
      __threadfence();                                         // cumulative over the CTA's reductions (bar.sync above)
      atomicAdd(args.grid_sync, 1);
      int seen;
      do {
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(args.grid_sync) : "memory");
        if (seen < static_cast<int>(gridDim.x)) __nanosleep(100);
      } while (seen < static_cast<int>(gridDim.x));
      __threadfence();
    }
    __syncthreads();
    const int nwarps = static_cast<int>(blockDim.x >> 5);
    const __nv_bfloat16* yo_rows = args.yacc ? nullptr : args.out0;       // gather mode: the rows every CTA just stored
    if (args.ln_gamma != nullptr) {
      // fused post-norm (hybrid_encoder.py:158-160): a warp per token, N <= 1024
      for (int t = blockIdx.x * nwarps + warp; t < args.T; t += gridDim.x * nwarps)
        ln_forward_token<kLnMaxI>(yo_rows, args.row_of_assignment, args.topk_gate, args.yacc, true, args.ln_res,
                                  args.ln_res_f32 != 0, args.ln_gamma, args.ln_beta, args.ln_eps, t, args.topk, args.N,
                                  lane, args.yout, args.ln_out_f32 != 0, args.ln_mean_rstd);
    } else if (args.yacc == nullptr) {
      for (int t = blockIdx.x * nwarps + warp; t < args.T; t += gridDim.x * nwarps)
        combine_token(yo_rows, args.row_of_assignment, args.topk_gate, true, t, args.topk, args.N, lane,
                      static_cast<__nv_bfloat16*>(args.yout));
    } else {
      const size_t n8 = static_cast<size_t>(args.T) * args.N / 8;
      const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
      __nv_bfloat16* yb = static_cast<__nv_bfloat16*>(args.yout);
      for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
        const float4 a = __ldcg(reinterpret_cast<const float4*>(args.yacc) + 2 * i);
        const float4 b = __ldcg(reinterpret_cast<const float4*>(args.yacc) + 2 * i + 1);
        const float o8[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        stg128(yb + 8 * i, pack8(o8));
      }
    }
  }
}

#undef FOR_EACH_TILE
}  // namespace b200moe

#include "gemm_host.inc"
