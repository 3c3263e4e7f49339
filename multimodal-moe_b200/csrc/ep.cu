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
