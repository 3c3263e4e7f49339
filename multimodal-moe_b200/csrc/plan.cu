// Capacity plan: priority scan -> per-expert offsets -> row maps.
//
// Priority of an assignment is (slot, token index): all first choices in token order, then all second choices.
// The router already counted assignments per (router block, slot, expert); plan_scan_kernel turns those counts into
// exclusive bases with warp scans (one warp per (slot, expert) sequence over the blocks, fixed order), clamps them
// to the capacity, lays the experts out back to back with each segment padded to `row_pad` rows, and reduces the
// loss statistics in a fixed order.  plan_assign_kernel then gives every assignment its row: base of its block +
// its rank among the same-expert assignments of the block (match_any + popc), which is exactly the slot-major /
// token-major order, so rows inside an expert are sorted by priority and "kept" == "row below the slot limit".
#include "common.cuh"

namespace b200moe {

constexpr int kScanThreads = 1024;

__global__ void __launch_bounds__(kScanThreads)
plan_scan_kernel(int32_t* __restrict__ block_counts, const float* __restrict__ block_stats, int nblk, int T, int k,
                 int E, int capacity, const int32_t* __restrict__ global_offset,
                 const int32_t* __restrict__ expert_order, int row_pad, int rows_cap, float aux_w, float z_w,
                 int32_t* __restrict__ expert_offsets, int32_t* __restrict__ meta, int32_t* __restrict__ tile_expert,
                 int32_t* __restrict__ tokens_per_expert, int32_t* __restrict__ kept_per_expert,
                 float* __restrict__ losses, int32_t* __restrict__ slot_limit, int32_t* __restrict__ token_of_row,
                 float* __restrict__ gate_of_row) {
  __shared__ int total_s[8 * 64];      // [k][E] assignments per (slot, expert) on this rank
  __shared__ int rowbase_s[8 * 64];    // [k][E] first row (relative to the expert segment) of slot s
  __shared__ int keptslot_s[8 * 64];
  __shared__ int kept_s[64], count_s[64], off_s[65], expert_at_s[64];
  __shared__ float P_s[65];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = kScanThreads / 32;
  const int kE = k * E;
  pdl_sync();

  // phase 1: exclusive scan over the router blocks for every (slot, expert) column, and the column sums of the loss
  // statistics.  Lanes run along the columns (one 128-byte line per block and warp: a lane-per-block layout made every
  // load and store touch 32 lines and the one SM's load/store unit, not memory latency, set the time: 30 us at 512
  // blocks), the 32 warps split the blocks into contiguous chunks: chunk sums of every column -> shared memory -> one
  // barrier -> prefix over the chunks -> second pass writes the bases.
  {
    extern __shared__ int scan_dyn[];
    int* chunk_s = scan_dyn;                                                  // [nwarps][kE]
    float* fchunk_s = reinterpret_cast<float*>(scan_dyn + nwarps * kE);       // [nwarps][E+1]
    const int ch = (nblk + nwarps - 1) / nwarps;
    const int b_lo = warp * ch, b_hi = (b_lo + ch < nblk) ? b_lo + ch : nblk;
    for (int q = lane; q < kE; q += 32) {
      int sum = 0;
      for (int b0 = b_lo; b0 < b_hi; b0 += 8) {
        int v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = (b0 + u < b_hi) ? block_counts[static_cast<size_t>(b0 + u) * kE + q] : 0;
#pragma unroll
        for (int u = 0; u < 8; ++u) sum += v[u];
      }
      chunk_s[warp * kE + q] = sum;
    }
    for (int e = lane; e <= E; e += 32) {
      float sum = 0.f;
      for (int b0 = b_lo; b0 < b_hi; b0 += 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = (b0 + u < b_hi) ? block_stats[static_cast<size_t>(b0 + u) * (E + 1) + e] : 0.f;
#pragma unroll
        for (int u = 0; u < 8; ++u) sum += v[u];
      }
      fchunk_s[warp * (E + 1) + e] = sum;
    }
    __syncthreads();
    for (int q = lane; q < kE; q += 32) {
      int run = 0, tot = 0;
      for (int w = 0; w < nwarps; ++w) {
        const int c = chunk_s[w * kE + q];
        run += (w < warp) ? c : 0;
        tot += c;
      }
      if (warp == 0) total_s[q] = tot;
      for (int b0 = b_lo; b0 < b_hi; b0 += 8) {
        int v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = (b0 + u < b_hi) ? block_counts[static_cast<size_t>(b0 + u) * kE + q] : 0;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          if (b0 + u < b_hi) block_counts[static_cast<size_t>(b0 + u) * kE + q] = run;
          run += v[u];
        }
      }
    }
    for (int e = threadIdx.x; e <= E; e += blockDim.x) {      // the chunks in a fixed order
      float sacc = 0.f;
      for (int w = 0; w < nwarps; ++w) sacc += fchunk_s[w * (E + 1) + e];
      P_s[e] = sacc;
    }
    __syncthreads();
  }

  // phase 2: per expert, slot bases against the capacity
  if (threadIdx.x < E) {
    const int e = threadIdx.x;
    int lower = 0, kept = 0, count = 0;
    for (int s = 0; s < k; ++s) {
      const int n = total_s[s * E + e];
      const int base = global_offset ? global_offset[s * E + e] : lower;  // earlier-priority assignments
      int room = capacity - base;
      room = room < 0 ? 0 : room;
      const int ks = n < room ? n : room;
      keptslot_s[s * E + e] = ks;
      rowbase_s[s * E + e] = kept;
      kept += ks;
      lower += n;
      count += n;
    }
    kept_s[e] = kept;
    count_s[e] = count;
    tokens_per_expert[e] = count;
    kept_per_expert[e] = kept;
    expert_at_s[expert_order ? expert_order[e] : e] = e;
  }
  __syncthreads();

  // phase 3: segment offsets in layout order, losses
  if (threadIdx.x == 0) {
    int off = 0, kept_total = 0, count_total = 0;
    for (int i = 0; i < E; ++i) {
      off_s[i] = off;
      const int e = expert_at_s[i];
      off += (kept_s[e] + row_pad - 1) / row_pad * row_pad;
      kept_total += kept_s[e];
      count_total += count_s[e];
    }
    off_s[E] = off;
    meta[0] = off;
    meta[1] = (off + kRowPad - 1) / kRowPad;
    meta[2] = count_total - kept_total;
    meta[3] = kept_total;
    meta[4] = capacity;
    meta[5] = meta[6] = meta[7] = 0;
    float lb = 0.f;
    const float invT = T > 0 ? 1.0f / static_cast<float>(T) : 0.f;
    for (int e = 0; e < E; ++e)
      lb += (static_cast<float>(count_s[e]) / static_cast<float>(static_cast<long long>(T) * k)) * (P_s[e] * invT);
    lb *= static_cast<float>(E);
    const float z = P_s[E] * invT;
    losses[0] = aux_w * lb + z_w * z;
    losses[1] = lb;
    losses[2] = z;
  }
  __syncthreads();
  for (int i = threadIdx.x; i <= E; i += blockDim.x) expert_offsets[i] = off_s[i];

  // phase 4: absolute first row and limit of every (slot, expert) for the assign kernel (slot_limit [2][k*E])
  for (int q = threadIdx.x; q < kE; q += blockDim.x) {
    const int e = q % E;
    const int pos = expert_order ? expert_order[e] : e;
    const int first = off_s[pos] + rowbase_s[q];
    slot_limit[q] = first + keptslot_s[q];
    slot_limit[kE + q] = first;
  }
  if (tile_expert) {
    const int ntiles_cap = rows_cap / kRowPad;
    const int ntiles = off_s[E] / kRowPad;
    for (int tIdx = threadIdx.x; tIdx < ntiles_cap; tIdx += blockDim.x) {
      int e = -1;
      if (tIdx < ntiles) {
        const int row = tIdx * kRowPad;
        int pos = 0;
        while (pos + 1 < E && off_s[pos + 1] <= row) ++pos;
        e = expert_at_s[pos];
      }
      tile_expert[tIdx] = e;
    }
  }
}

__global__ void __launch_bounds__(kRouterTokPerBlock)
plan_assign_kernel(const int32_t* __restrict__ topk_idx, const float* __restrict__ topk_gate,
                   const int32_t* __restrict__ block_base, const int32_t* __restrict__ slot_limit, int T, int k,
                   int E, int32_t* __restrict__ row_of_assignment, int32_t* __restrict__ token_of_row,
                   float* __restrict__ gate_of_row, const int32_t* __restrict__ expert_offsets,
                   const int32_t* __restrict__ kept_per_expert, const int32_t* __restrict__ expert_order,
                   int rows_cap) {
  __shared__ int wc[kRouterTokPerBlock / 32][64];
  __shared__ int off_s[65], end_s[64];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = t < T;
  pdl_sync();
  // pad rows (rows no assignment lands on) are marked here, a slice per block: no memset launches before the plan
  for (int i = threadIdx.x; i <= E; i += blockDim.x) off_s[i] = expert_offsets[i];
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    const int pos = expert_order ? expert_order[e] : e;
    end_s[pos] = expert_offsets[pos] + kept_per_expert[e];
  }
  __syncthreads();
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < rows_cap; r += gridDim.x * blockDim.x) {
    bool pad = r >= off_s[E];
    if (!pad) {
      int lo = 0, hi = E;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (off_s[mid] <= r) lo = mid; else hi = mid;
      }
      while (lo + 1 < E && off_s[lo + 1] <= r) ++lo;     // skip empty segments that start at the same row
      pad = r >= end_s[lo];
    }
    if (pad) { token_of_row[r] = -1; gate_of_row[r] = 0.f; }
  }
  for (int s = 0; s < k; ++s) {
    for (int i = threadIdx.x; i < (kRouterTokPerBlock / 32) * 64; i += blockDim.x) (&wc[0][0])[i] = 0;
    __syncthreads();
    const int e = live ? topk_idx[static_cast<size_t>(t) * k + s] : (0x40000000 + lane);
    const unsigned same = __match_any_sync(0xffffffffu, e);
    const int rank_in_warp = __popc(same & ((1u << lane) - 1u));
    if (live && rank_in_warp == 0) wc[warp][e] = __popc(same);
    __syncthreads();
    if (live) {
      int rank = rank_in_warp;
      for (int w = 0; w < warp; ++w) rank += wc[w][e];
      const int q = s * E + e;
      const int row = slot_limit[k * E + q] + block_base[static_cast<size_t>(blockIdx.x) * k * E + q] + rank;
      const bool keep = row < slot_limit[q];
      row_of_assignment[static_cast<size_t>(t) * k + s] = keep ? row : -1;
      if (keep) {
        token_of_row[row] = t;
        gate_of_row[row] = topk_gate[static_cast<size_t>(t) * k + s];
      }
    }
    __syncthreads();
  }
}

// Small problems (<= kFusedMaxBlocks router blocks): one launch.  Every block redoes the (tiny) scan for itself --
// totals and its own exclusive base from the nblk x k x E counts -- and the capacity / offset arithmetic in shared
// memory, then assigns the rows of its 128 tokens; block 0 also writes the plan's global outputs and the losses, and
// the blocks share the initialisation of the pad rows (rows no assignment lands on), so no second kernel and no
// single-block stage sits on the critical path of a latency-bound step.
constexpr int kFusedMaxBlocks = 64;

__global__ void __launch_bounds__(kRouterTokPerBlock)
plan_fused_kernel(const int32_t* __restrict__ topk_idx, const float* __restrict__ topk_gate,
                  const int32_t* __restrict__ block_counts, const float* __restrict__ block_stats, int nblk, int T,
                  int k, int E, int capacity, const int32_t* __restrict__ global_offset,
                  const int32_t* __restrict__ expert_order, int row_pad, int rows_cap, float aux_w, float z_w,
                  int32_t* __restrict__ expert_offsets, int32_t* __restrict__ meta, int32_t* __restrict__ tile_expert,
                  int32_t* __restrict__ tokens_per_expert, int32_t* __restrict__ kept_per_expert,
                  float* __restrict__ losses, int32_t* __restrict__ slot_limit,
                  int32_t* __restrict__ row_of_assignment, int32_t* __restrict__ token_of_row,
                  float* __restrict__ gate_of_row) {
  __shared__ int total_s[8 * 64], base_s[8 * 64], first_s[8 * 64], keptslot_s[8 * 64];
  __shared__ int kept_s[64], count_s[64], off_s[65], expert_at_s[64];
  __shared__ float P_s[65];
  __shared__ int wc[kRouterTokPerBlock / 32][64];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = kRouterTokPerBlock / 32;
  const int kE = k * E;
  const bool lead = blockIdx.x == 0;
  pdl_sync();

  // totals and this block's exclusive base of every (slot, expert) column: a warp per column, the lanes over the (at most
  // kFusedMaxBlocks = 64) router blocks, so every load of the scan is in flight at once (a thread per column walked the
  // blocks one dependent load after the other: the longest latency chain of the kernel at 3200 tokens)
  for (int q = warp; q < kE; q += nwarps) {
    const int c0 = lane < nblk ? block_counts[static_cast<size_t>(lane) * kE + q] : 0;
    const int c1 = lane + 32 < nblk ? block_counts[static_cast<size_t>(lane + 32) * kE + q] : 0;
    const int me = static_cast<int>(blockIdx.x);
    int tot = c0 + c1;
    int base = (lane < me ? c0 : 0) + (lane + 32 < me ? c1 : 0);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      tot += __shfl_xor_sync(0xffffffffu, tot, o);
      base += __shfl_xor_sync(0xffffffffu, base, o);
    }
    if (lane == 0) { total_s[q] = tot; base_s[q] = base; }
  }
  if (lead) {   // loss statistics: one warp per column, lanes stride the blocks, fixed order
    for (int e = warp; e <= E; e += nwarps) {
      float s = 0.f;
      for (int b0 = 0; b0 < nblk; b0 += 256) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int b = b0 + u * 32 + lane;
          v[u] = (b < nblk) ? block_stats[static_cast<size_t>(b) * (E + 1) + e] : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) s += v[u];
      }
      s = warp_sum(s);
      if (lane == 0) P_s[e] = s;
    }
  }
  __syncthreads();
  if (threadIdx.x < E) {
    const int e = threadIdx.x;
    int lower = 0, kept = 0, count = 0;
    for (int s = 0; s < k; ++s) {
      const int n = total_s[s * E + e];
      const int base = global_offset ? global_offset[s * E + e] : lower;
      int room = capacity - base;
      room = room < 0 ? 0 : room;
      const int ks = n < room ? n : room;
      keptslot_s[s * E + e] = ks;
      first_s[s * E + e] = kept;          // relative to the expert segment until the offsets are known
      kept += ks;
      lower += n;
      count += n;
    }
    kept_s[e] = kept;
    count_s[e] = count;
    expert_at_s[expert_order ? expert_order[e] : e] = e;
    if (lead) {
      tokens_per_expert[e] = count;
      kept_per_expert[e] = kept;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int off = 0, kept_total = 0, count_total = 0;
    for (int i = 0; i < E; ++i) {
      off_s[i] = off;
      const int e = expert_at_s[i];
      off += (kept_s[e] + row_pad - 1) / row_pad * row_pad;
      kept_total += kept_s[e];
      count_total += count_s[e];
    }
    off_s[E] = off;
    if (lead) {
      meta[0] = off;
      meta[1] = (off + kRowPad - 1) / kRowPad;
      meta[2] = count_total - kept_total;
      meta[3] = kept_total;
      meta[4] = capacity;
      meta[5] = meta[6] = meta[7] = 0;
      float lb = 0.f;
      const float invT = T > 0 ? 1.0f / static_cast<float>(T) : 0.f;
      for (int e = 0; e < E; ++e)
        lb += (static_cast<float>(count_s[e]) / static_cast<float>(static_cast<long long>(T) * k)) * (P_s[e] * invT);
      lb *= static_cast<float>(E);
      const float z = P_s[E] * invT;
      losses[0] = aux_w * lb + z_w * z;
      losses[1] = lb;
      losses[2] = z;
    }
  }
  __syncthreads();
  for (int q = threadIdx.x; q < kE; q += blockDim.x) {
    const int e = q % E;
    const int pos = expert_order ? expert_order[e] : e;
    const int first = off_s[pos] + first_s[q];
    first_s[q] = first;
    if (lead) {
      slot_limit[q] = first + keptslot_s[q];
      slot_limit[kE + q] = first;
    }
  }
  if (lead)
    for (int i = threadIdx.x; i <= E; i += blockDim.x) expert_offsets[i] = off_s[i];
  // pad rows and tile -> expert map, shared between the blocks
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < rows_cap; r += gridDim.x * blockDim.x) {
    int pos = -1;
    if (r < off_s[E]) {
      pos = 0;
      while (pos + 1 < E && off_s[pos + 1] <= r) ++pos;
    }
    const int e = pos >= 0 ? expert_at_s[pos] : -1;
    if (pos < 0 || r >= off_s[pos] + kept_s[e]) { token_of_row[r] = -1; gate_of_row[r] = 0.f; }
    if (tile_expert && (r % kRowPad) == 0) tile_expert[r / kRowPad] = e;
  }
  __syncthreads();

  // rows of this block's assignments (as plan_assign_kernel)
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = t < T;
  for (int s = 0; s < k; ++s) {
    for (int i = threadIdx.x; i < (kRouterTokPerBlock / 32) * 64; i += blockDim.x) (&wc[0][0])[i] = 0;
    __syncthreads();
    const int e = live ? topk_idx[static_cast<size_t>(t) * k + s] : (0x40000000 + lane);
    const unsigned same = __match_any_sync(0xffffffffu, e);
    const int rank_in_warp = __popc(same & ((1u << lane) - 1u));
    if (live && rank_in_warp == 0) wc[warp][e] = __popc(same);
    __syncthreads();
    if (live) {
      int rank = rank_in_warp;
      for (int w = 0; w < warp; ++w) rank += wc[w][e];
      const int q = s * E + e;
      const int row = first_s[q] + base_s[q] + rank;
      const bool keep = row < first_s[q] + keptslot_s[q];
      row_of_assignment[static_cast<size_t>(t) * k + s] = keep ? row : -1;
      if (keep) {
        token_of_row[row] = t;
        gate_of_row[row] = topk_gate[static_cast<size_t>(t) * k + s];
      }
    }
    __syncthreads();
  }
}

}  // namespace b200moe

extern "C" int b200moe_plan(const int32_t* topk_idx, const float* topk_gate, int32_t* block_counts,
                            const float* block_stats, int T, int k, int E, int capacity,
                            const int32_t* global_offset, const int32_t* expert_order, int row_pad, int rows_cap,
                            float aux_loss_weight, float z_loss_weight, int32_t* expert_offsets, int32_t* meta,
                            int32_t* tile_expert, int32_t* tokens_per_expert, int32_t* kept_per_expert,
                            float* losses, int32_t* slot_limit, int32_t* row_of_assignment, int32_t* token_of_row,
                            float* gate_of_row, void* stream) {
  using namespace b200moe;
  B200MOE_REQUIRE(T > 0, B200MOE_EINVAL, "plan: T=%d must be positive", T);
  B200MOE_REQUIRE(E >= 1 && E <= 64 && k >= 1 && k <= 8 && k <= E, B200MOE_EINVAL, "plan: bad E=%d / k=%d", E, k);
  B200MOE_REQUIRE(row_pad == 1 || row_pad == kRowPad || row_pad == 2 * kRowPad, B200MOE_EINVAL,
                  "plan: row_pad must be 1, 128 or 256");
  B200MOE_REQUIRE(capacity >= 0, B200MOE_EINVAL, "plan: negative capacity");
  B200MOE_REQUIRE(rows_cap > 0 && rows_cap % kRowPad == 0, B200MOE_EINVAL, "plan: rows_cap must be a multiple of 128");
  if (int rc = b200moe_check_device()) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int tb = router_block_tokens(T);
  const int nblk = (T + tb - 1) / tb;
  if (nblk <= kFusedMaxBlocks) {
    launch_k(plan_fused_kernel, nblk, tb, 0, st, topk_idx, topk_gate, static_cast<const int32_t*>(block_counts),
             block_stats, nblk, T, k, E, capacity, global_offset, expert_order, row_pad, rows_cap, aux_loss_weight,
             z_loss_weight, expert_offsets, meta, tile_expert, tokens_per_expert, kept_per_expert, losses, slot_limit,
             row_of_assignment, token_of_row, gate_of_row);
    return check_launch("plan_fused");
  }
  const size_t scan_smem = sizeof(int) * (kScanThreads / 32) * (static_cast<size_t>(k) * E + E + 1);
  static DeviceOnce scan_once;
  const int dev = current_device();
  if (!scan_once.done(dev)) {
    cudaError_t e = cudaFuncSetAttribute(plan_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * (8 * 64 + 65) * 4);
    B200MOE_REQUIRE(e == cudaSuccess, B200MOE_ECUDA, "plan: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    scan_once.set(dev);
  }
  launch_k(plan_scan_kernel, 1, kScanThreads, scan_smem, st, block_counts, block_stats, nblk, T, k, E, capacity, global_offset,
           expert_order, row_pad, rows_cap, aux_loss_weight, z_loss_weight, expert_offsets, meta, tile_expert,
           tokens_per_expert, kept_per_expert, losses, slot_limit, token_of_row, gate_of_row);
  if (int rc = check_launch("plan_scan")) return rc;
  launch_k(plan_assign_kernel, nblk, tb, 0, st, topk_idx, topk_gate, block_counts, slot_limit, T, k, E,
           row_of_assignment, token_of_row, gate_of_row, static_cast<const int32_t*>(expert_offsets),
           static_cast<const int32_t*>(kept_per_expert), expert_order, rows_cap);
  return check_launch("plan_assign");
}
