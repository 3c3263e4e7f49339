// Expert-parallel receive side.  After the token all-to-all, rows sit back to back as segments
// (source rank major, local expert minor; any order is accepted).  This kernel lays them out as 128-row-padded expert segments for the
// grouped GEMMs and emits the row map used to gather them in and to scatter the results back out.
#include "common.cuh"

namespace b200moe {

__global__ void __launch_bounds__(256)
ep_recv_plan_kernel(const int32_t* __restrict__ recv_counts, const int32_t* __restrict__ seg_expert, int n_seg,
                    int E_local, int rows_cap, int32_t* __restrict__ expert_offsets, int32_t* __restrict__ meta,
                    int32_t* __restrict__ tile_expert, int32_t* __restrict__ src_of_row) {
  __shared__ int off_s[65];        // padded offsets per local expert
  __shared__ int cnt_s[64];        // rows per local expert
  __shared__ int seg_src_s[1024];  // first row of segment s in the receive buffer
  __shared__ int seg_dst_s[1024];  // first row of segment s in the padded layout
  if (threadIdx.x == 0) {
    for (int e = 0; e < E_local; ++e) cnt_s[e] = 0;
    int src = 0;
    for (int s = 0; s < n_seg; ++s) {
      seg_src_s[s] = src;
      seg_dst_s[s] = cnt_s[seg_expert[s]];          // position inside its expert, in arrival order
      cnt_s[seg_expert[s]] += recv_counts[s];
      src += recv_counts[s];
    }
    int off = 0;
    for (int e = 0; e < E_local; ++e) {
      off_s[e] = off;
      off += (cnt_s[e] + kRowPad - 1) / kRowPad * kRowPad;
    }
    off_s[E_local] = off;
    for (int s = 0; s < n_seg; ++s) seg_dst_s[s] += off_s[seg_expert[s]];
    meta[0] = off;
    meta[1] = off / kRowPad;
    meta[2] = 0;
    meta[3] = src;
    meta[4] = meta[5] = meta[6] = meta[7] = 0;
  }
  __syncthreads();
  for (int i = threadIdx.x; i <= E_local; i += blockDim.x) expert_offsets[i] = off_s[i];
  const int R = off_s[E_local];
  for (int r = threadIdx.x; r < rows_cap; r += blockDim.x) src_of_row[r] = -1;
  __syncthreads();
  for (int s = 0; s < n_seg; ++s) {
    const int n = recv_counts[s], src = seg_src_s[s], dst = seg_dst_s[s];
    for (int j = threadIdx.x; j < n; j += blockDim.x) src_of_row[dst + j] = src + j;
  }
  for (int t = threadIdx.x; t < rows_cap / kRowPad; t += blockDim.x) {
    int e = -1;
    if (t * kRowPad < R) {
      e = 0;
      while (e + 1 < E_local && off_s[e + 1] <= t * kRowPad) ++e;
    }
    tile_expert[t] = e;
  }
}

// hist[s,e] = sum over the router blocks of block_counts[b,s,e]; out[k*E] = T.  The per-rank message of the histogram
// all-gather, built in one launch.
__global__ void __launch_bounds__(256)
ep_hist_kernel(const int32_t* __restrict__ block_counts, int nblk, int kE, int T, int32_t* __restrict__ out) {
  for (int q = threadIdx.x; q < kE; q += blockDim.x) {
    int s = 0;
    for (int b = 0; b < nblk; ++b) s += block_counts[static_cast<size_t>(b) * kE + q];
    out[q] = s;
  }
  if (threadIdx.x == 0) out[kE] = T;
}

// Global routing tables from the gathered histograms (the device twin of b200moe.ep.route_tables, A.5 under EP):
//   capacity = ceil(cf * T_global * k / E) in double, evaluated in the same order as the host code
//   base[w,s,e] = assignments of earlier priority: all ranks' lower slots + lower ranks' same slot
//   kept[w,e]   = sum_s min(hist[w,s,e], max(capacity - base[w,s,e], 0))
// Outputs: goff [k,E] = base[rank]; kept_out [W*E] then [W*E] = capacity, [W*E+1] = T_global (copied to the host for
// the all-to-all split sizes); recv_counts [C][W][ELc] = kept[w, rank*EL + c*ELc + j] for ep_recv_plan.
__global__ void __launch_bounds__(64)
ep_route_tables_kernel(const int32_t* __restrict__ allh, int W, int k, int E, int rank, double cf, int ELc,
                       int32_t* __restrict__ goff, int32_t* __restrict__ kept_out, int32_t* __restrict__ recv_counts) {
  const int stride = k * E + 1;
  long long tg = 0;
  for (int w = 0; w < W; ++w) tg += allh[static_cast<size_t>(w) * stride + k * E];
  long long cap;
  if (cf < 0.0) cap = tg * k;
  else cap = static_cast<long long>(ceil(cf * static_cast<double>(tg) * static_cast<double>(k) / static_cast<double>(E)));
  const int e = threadIdx.x;
  if (e < E) {
    long long lower = 0;                       // all ranks, earlier slot levels
    long long kept_w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int s = 0; s < k; ++s) {
      long long before = 0;                    // same slot level, lower ranks
      for (int w = 0; w < W; ++w) {
        const long long h = allh[static_cast<size_t>(w) * stride + s * E + e];
        const long long base = lower + before;
        long long room = cap - base;
        room = room < 0 ? 0 : room;
        kept_w[w] += h < room ? h : room;
        if (w == rank) goff[s * E + e] = static_cast<int32_t>(base);
        before += h;
      }
      lower += before;
    }
    const int EL = E / W, C = EL / ELc;
    for (int w = 0; w < W; ++w) {
      kept_out[w * E + e] = static_cast<int32_t>(kept_w[w]);
      if (e / EL == rank) {
        const int el = e - rank * EL, c = el / ELc, j = el - c * ELc;
        recv_counts[(c * W + w) * ELc + j] = static_cast<int32_t>(kept_w[w]);
      }
    }
    (void)C;
  }
  if (threadIdx.x == 0) {
    kept_out[W * E] = static_cast<int32_t>(cap > 0x7fffffffLL ? 0x7fffffffLL : cap);
    kept_out[W * E + 1] = static_cast<int32_t>(tg);
  }
}

}  // namespace b200moe

extern "C" int b200moe_ep_recv_plan(const int32_t* recv_counts, const int32_t* seg_expert, int n_seg, int E_local,
                                    int rows_cap, int32_t* expert_offsets, int32_t* meta, int32_t* tile_expert,
                                    int32_t* src_of_row, void* stream) {
  using namespace b200moe;
  B200MOE_REQUIRE(E_local >= 1 && E_local <= 64 && n_seg >= 1 && n_seg <= 1024, B200MOE_EINVAL, "ep_recv_plan: bad E_local=%d / n_seg=%d", E_local, n_seg);
  B200MOE_REQUIRE(rows_cap > 0 && rows_cap % kRowPad == 0, B200MOE_EINVAL, "ep_recv_plan: rows_cap must be a multiple of 128");
  if (int rc = b200moe_check_device()) return rc;
  ep_recv_plan_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(recv_counts, seg_expert, n_seg, E_local,
                                                                        rows_cap, expert_offsets, meta, tile_expert,
                                                                        src_of_row);
  return check_launch("ep_recv_plan");
}

extern "C" int b200moe_ep_hist(const int32_t* block_counts, int nblk, int k, int E, int T, int32_t* out, void* stream) {
  using namespace b200moe;
  B200MOE_REQUIRE(nblk >= 1 && E >= 1 && E <= 64 && k >= 1 && k <= 8, B200MOE_EINVAL, "ep_hist: bad shape");
  if (int rc = b200moe_check_device()) return rc;
  ep_hist_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(block_counts, nblk, k * E, T, out);
  return check_launch("ep_hist");
}

extern "C" int b200moe_ep_route_tables(const int32_t* all_hist, int W, int k, int E, int rank, double capacity_factor,
                                       int experts_per_chunk, int32_t* global_offset, int32_t* kept,
                                       int32_t* recv_counts, void* stream) {
  using namespace b200moe;
  B200MOE_REQUIRE(W >= 1 && W <= 8 && E >= 1 && E <= 64 && E % W == 0 && k >= 1 && k <= 8 && rank >= 0 && rank < W,
                  B200MOE_EINVAL, "ep_route_tables: bad W=%d / E=%d / k=%d / rank=%d", W, E, k, rank);
  B200MOE_REQUIRE(experts_per_chunk >= 1 && (E / W) % experts_per_chunk == 0, B200MOE_EINVAL,
                  "ep_route_tables: experts_per_chunk=%d does not divide E/W", experts_per_chunk);
  if (int rc = b200moe_check_device()) return rc;
  ep_route_tables_kernel<<<1, 64, 0, static_cast<cudaStream_t>(stream)>>>(all_hist, W, k, E, rank, capacity_factor,
                                                                          experts_per_chunk, global_offset, kept,
                                                                          recv_counts);
  return check_launch("ep_route_tables");
}
