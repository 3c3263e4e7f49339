// Expert-parallel dispatch / combine over peer memory: NVLink loads and stores from inside the
// row-movement kernels instead of an all-to-all between them.
//
// Every rank owns an arena of four row buffers [rows_cap, d] bf16 -- xp (expert inputs), yo (expert outputs), dyp
// (their gradients), dxp (input gradients) -- in the padded expert-major layout the grouped GEMMs read, mapped into
// the address space of every other rank (symmetric memory).  After the histogram all-gather every rank knows
// kept[W,E], hence the layout of EVERY rank's buffers, so
//   * dispatch  : the home rank gathers x[token] and stores the row straight into the expert rank's xp (one kernel:
//                 permute + transfer; no send buffer, no receive buffer, no second permute);
//   * combine   : the home rank loads the k output rows of a token straight out of the expert ranks' yo;
//   * backward  : dyp rows are stored into the expert rank (the dot product for the gate gradient loads the yo row
//                 over NVLink), dxp rows are loaded by the home rank in unpermute_bwd.
// Between a producer and its remote consumers sits one device-side barrier over the ranks (the host side uses the
// symmetric-memory signal pads); there is no host synchronisation and no split-size exchange anywhere in the pass.
#include "common.cuh"

namespace b200moe {

constexpr int kP2pWarps = 8;
constexpr int kP2pThreads = kP2pWarps * 32;

// kept [W,E] (+ capacity, T_global as written by ep_route_tables) -> the layout tables.
//   dest_off[e]       : first row, inside the buffer of rank e / EL, of the rows THIS rank sends to expert e
//                       (segment of expert e on its rank + rows of the lower ranks: rank-major inside an expert)
//   expert_offsets[j] : this rank's own padded segments (local expert j), meta / tile_expert as b200moe_plan writes them
//   meta[5] = 1 if some rank's rows do not fit rows_cap (cannot happen while every rank stays within the token count the
//   arena was sized for -- the host enforces that; rows beyond rows_cap are never stored, the flag is the evidence)
__global__ void __launch_bounds__(64)
ep_p2p_tables_kernel(const int32_t* __restrict__ kept, int W, int E, int rank, int rows_cap, int row_pad,
                     int32_t* __restrict__ dest_off, int32_t* __restrict__ expert_offsets, int32_t* __restrict__ meta,
                     int32_t* __restrict__ tile_expert) {
  __shared__ int tot_s[64], off_s[65], real_s[1], bad_s[1];
  const int EL = E / W;
  const int e = threadIdx.x;
  if (e == 0) { bad_s[0] = 0; real_s[0] = 0; }
  if (e < E) {
    int tot = 0;
    for (int w = 0; w < W; ++w) tot += kept[w * E + e];
    tot_s[e] = tot;
  }
  __syncthreads();
  if (e < E) {
    const int w = e / EL;
    int off = 0;
    for (int j = w * EL; j < e; ++j) off += (tot_s[j] + row_pad - 1) / row_pad * row_pad;
    int before = 0;
    for (int s = 0; s < rank; ++s) before += kept[s * E + e];
    dest_off[e] = off + before;
    if (w == rank) off_s[e - w * EL] = off;
    if (e == w * EL + EL - 1) {       // last expert of rank w: total rows of that rank
      const int total = off + (tot_s[e] + row_pad - 1) / row_pad * row_pad;
      if (total > rows_cap) bad_s[0] = 1;
      if (w == rank) off_s[EL] = total;
    }
    if (w == rank) atomicAdd(&real_s[0], tot_s[e]);
  }
  __syncthreads();
  for (int j = threadIdx.x; j <= EL; j += blockDim.x) expert_offsets[j] = off_s[j];
  if (threadIdx.x == 0) {
    const int R = off_s[EL] <= rows_cap ? off_s[EL] : 0;
    meta[0] = R;
    meta[1] = R / kRowPad;
    meta[2] = 0;
    meta[3] = real_s[0];
    meta[4] = 0;
    meta[5] = bad_s[0];
    meta[6] = meta[7] = 0;
  }
  const int R = off_s[EL] <= rows_cap ? off_s[EL] : 0;
  for (int t = threadIdx.x; t < rows_cap / kRowPad; t += blockDim.x) {
    int j = -1;
    if (t * kRowPad < R) {
      j = 0;
      while (j + 1 < EL && off_s[j + 1] <= t * kRowPad) ++j;
    }
    tile_expert[t] = j;
  }
}

// Zero the pad rows of up to two local buffers (xp and dyp): rows [off_j + n_j, off_{j+1}) of every local expert.
__global__ void __launch_bounds__(kP2pThreads)
ep_p2p_zero_pads_kernel(const int32_t* __restrict__ kept, const int32_t* __restrict__ expert_offsets, int W, int E,
                        int rank, int d, __nv_bfloat16* __restrict__ a, __nv_bfloat16* __restrict__ b) {
  const int EL = E / W;
  const int j = blockIdx.y;
  int n = 0;
  for (int w = 0; w < W; ++w) n += kept[w * E + rank * EL + j];
  const int lo = expert_offsets[j] + n, hi = expert_offsets[j + 1];
  const int nchunk = d >> 3, lane = threadIdx.x & 31;
  for (int r = lo + blockIdx.x * kP2pWarps + (threadIdx.x >> 5); r < hi; r += gridDim.x * kP2pWarps) {
    for (int c = lane; c < nchunk; c += 32) {
      if (a) stg128(a + static_cast<size_t>(r) * d + c * 8, make_uint4(0, 0, 0, 0));
      if (b) stg128(b + static_cast<size_t>(r) * d + c * 8, make_uint4(0, 0, 0, 0));
    }
  }
}

__device__ __forceinline__ int expert_of_send_row(const int* off_s, int E, int r) {
  int lo = 0, hi = E;                  // largest e with off[e] <= r  (offsets are non-decreasing, off[E] > r)
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (off_s[mid] <= r) lo = mid; else hi = mid;
  }
  return lo;
}

// Dispatch: send row r (send layout of b200moe_plan with row_pad = 1: experts back to back, priority order inside an
// expert) -> row dest_off[e] + (r - send_off[e]) of rank e / EL.  Also records dest_of_row[r] for the later passes.
__global__ void __launch_bounds__(kP2pThreads)
ep_p2p_dispatch_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ token_of_row,
                       const int32_t* __restrict__ send_off, const int32_t* __restrict__ send_meta,
                       const int32_t* __restrict__ dest_off, PeerRows pr, int E, int EL, int d,
                       int32_t* __restrict__ dest_of_row) {
  __shared__ int off_s[65], dst_s[64];
  for (int i = threadIdx.x; i <= E; i += blockDim.x) off_s[i] = send_off[i];
  for (int i = threadIdx.x; i < E; i += blockDim.x) dst_s[i] = dest_off[i];
  __syncthreads();
  const int R = send_meta[0];
  const int nchunk = d >> 3, lane = threadIdx.x & 31;
  for (int r = blockIdx.x * kP2pWarps + (threadIdx.x >> 5); r < R; r += gridDim.x * kP2pWarps) {
    const int e = expert_of_send_row(off_s, E, r);
    const int row = dst_s[e] + (r - off_s[e]);
    const int w = e / EL;
    const bool ok = row < pr.rows_cap;
    const int dest = ok ? w * pr.rows_cap + row : -1;
    if (lane == 0) dest_of_row[r] = dest;
    const int t = token_of_row[r];
    if (!ok || t < 0) continue;
    const __nv_bfloat16* src = x + static_cast<size_t>(t) * d;
    __nv_bfloat16* dst = peer_row(pr, dest, d);
    for (int c = lane; c < nchunk; c += 32) stg128(dst + c * 8, ldg128(src + c * 8));
  }
}

// y[t] = sum_s gate[t,s] * yo_{rank(t,s)}[row(t,s)]   (slot order, fp32): the rows come over NVLink
__global__ void __launch_bounds__(kP2pThreads)
ep_p2p_combine_kernel(PeerRows pr, const int32_t* __restrict__ row_of_assignment,
                      const float* __restrict__ topk_gate, int T, int k, int d, __nv_bfloat16* __restrict__ y) {
  const int nchunk = d >> 3, lane = threadIdx.x & 31;
  for (int t = blockIdx.x * kP2pWarps + (threadIdx.x >> 5); t < T; t += gridDim.x * kP2pWarps) {
    const __nv_bfloat16* src[8];
    float g[8];
    for (int s = 0; s < k; ++s) {
      const int r = row_of_assignment[static_cast<size_t>(t) * k + s];
      const int dest = r >= 0 ? pr.dest_of_row[r] : -1;
      src[s] = dest >= 0 ? peer_row(pr, dest, d) : nullptr;
      g[s] = topk_gate[static_cast<size_t>(t) * k + s];
    }
    for (int c = lane; c < nchunk; c += 32) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int s = 0; s < k; ++s) {
        if (!src[s]) continue;
        float v[8];
        unpack8(ldg128(src[s] + c * 8), v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = combine_add(acc[j], combine_term(g[s], v[j]));
      }
      stg128(y + static_cast<size_t>(t) * d + c * 8, pack8(acc));
    }
  }
}

// Backward of combine at the home rank: dyp_{rank}[row] = gate * dy[token] (stored into the expert rank),
// dgate_row[r] = <dy[token], yo_{rank}[row]> (yo row loaded from the expert rank).
__global__ void __launch_bounds__(kP2pThreads)
ep_p2p_combine_bwd_kernel(const __nv_bfloat16* __restrict__ dy, PeerRows yo, PeerRows dyp,
                          const int32_t* __restrict__ token_of_row, const float* __restrict__ gate_of_row,
                          const int32_t* __restrict__ send_meta, int d, float* __restrict__ dgate_row) {
  const int R = send_meta[0];
  const int nchunk = d >> 3, lane = threadIdx.x & 31;
  for (int r = blockIdx.x * kP2pWarps + (threadIdx.x >> 5); r < R; r += gridDim.x * kP2pWarps) {
    const int t = token_of_row[r];
    const int dest = yo.dest_of_row[r];
    if (t < 0 || dest < 0) {
      if (lane == 0) dgate_row[r] = 0.f;
      continue;
    }
    const float g = gate_of_row[r];
    const __nv_bfloat16* yrow = peer_row(yo, dest, d);
    __nv_bfloat16* drow = peer_row(dyp, dest, d);
    float dot = 0.f;
    for (int c = lane; c < nchunk; c += 32) {
      float a[8], b[8], o[8];
      unpack8(ldg128(dy + static_cast<size_t>(t) * d + c * 8), a);
      unpack8(ldg128(yrow + c * 8), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        o[j] = g * a[j];
        dot = fmaf(a[j], b[j], dot);
      }
      stg128(drow + c * 8, pack8(o));
    }
    dot = warp_sum(dot);
    if (lane == 0) dgate_row[r] = dot;
  }
}

// out[r] = peers' buffer[dest_of_row[r]]: the return trip of the input gradients as a plain row copy.  No shared memory
// and few registers, so its blocks co-reside with the persistent weight-gradient GEMM CTAs: the NVLink transfer runs
// under the tensor-bound kernels instead of after them.
__global__ void __launch_bounds__(kP2pThreads)
ep_p2p_pull_rows_kernel(PeerRows pr, const int32_t* __restrict__ send_meta, int d, __nv_bfloat16* __restrict__ out) {
  const int R = send_meta[0];
  const int nchunk = d >> 3, lane = threadIdx.x & 31;
  for (int r = blockIdx.x * kP2pWarps + (threadIdx.x >> 5); r < R; r += gridDim.x * kP2pWarps) {
    const int dest = pr.dest_of_row[r];
    __nv_bfloat16* dst = out + static_cast<size_t>(r) * d;
    if (dest < 0) {
      for (int c = lane; c < nchunk; c += 32) stg128(dst + c * 8, make_uint4(0, 0, 0, 0));
      continue;
    }
    const __nv_bfloat16* src = peer_row(pr, dest, d);
    for (int c = lane; c < nchunk; c += 32) stg128(dst + c * 8, ldg128(src + c * 8));
  }
}

static int p2p_grid(int rows) {
  int g = (rows + kP2pWarps - 1) / kP2pWarps;
  const int mx = num_sms() * 8;
  return g < 1 ? 1 : (g > mx ? mx : g);
}

}  // namespace b200moe

using namespace b200moe;

extern "C" int b200moe_ep_p2p_tables(const int32_t* kept, int W, int E, int rank, int rows_cap, int row_pad,
                                     int32_t* dest_off, int32_t* expert_offsets, int32_t* meta, int32_t* tile_expert,
                                     void* stream) {
  B200MOE_REQUIRE(row_pad == kRowPad || row_pad == 2 * kRowPad, B200MOE_EINVAL, "ep_p2p_tables: row_pad must be 128 or 256");
  B200MOE_REQUIRE(W >= 1 && W <= 8 && E >= 1 && E <= 64 && E % W == 0 && rank >= 0 && rank < W, B200MOE_EINVAL,
                  "ep_p2p_tables: bad world=%d / num_experts=%d / rank=%d", W, E, rank);
  B200MOE_REQUIRE(rows_cap > 0 && rows_cap % kRowPad == 0, B200MOE_EINVAL, "ep_p2p_tables: rows_cap must be a multiple of 128");
  if (int rc = b200moe_check_device()) return rc;
  ep_p2p_tables_kernel<<<1, 64, 0, static_cast<cudaStream_t>(stream)>>>(kept, W, E, rank, rows_cap, row_pad, dest_off,
                                                                        expert_offsets, meta, tile_expert);
  return check_launch("ep_p2p_tables");
}

extern "C" int b200moe_ep_p2p_zero_pads(const int32_t* kept, const int32_t* expert_offsets, int W, int E, int rank,
                                        int d, void* a, void* b, void* stream) {
  B200MOE_REQUIRE(W >= 1 && E % W == 0 && d % 8 == 0, B200MOE_EINVAL, "ep_p2p_zero_pads: bad arguments");
  if (int rc = b200moe_check_device()) return rc;
  dim3 grid(2, E / W);
  ep_p2p_zero_pads_kernel<<<grid, kP2pThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      kept, expert_offsets, W, E, rank, d, static_cast<__nv_bfloat16*>(a), static_cast<__nv_bfloat16*>(b));
  return check_launch("ep_p2p_zero_pads");
}

extern "C" int b200moe_ep_p2p_dispatch(const void* x, const int32_t* token_of_row, const int32_t* send_offsets,
                                       const int32_t* send_meta, const int32_t* dest_off, const void* const* peers,
                                       int64_t buf_offset_bytes, int rows_cap, int send_rows_cap, int E, int E_local,
                                       int d, int32_t* dest_of_row, void* stream) {
  B200MOE_REQUIRE(x && peers && dest_of_row && d % 8 == 0 && E >= 1 && E <= 64 && E_local >= 1, B200MOE_EINVAL,
                  "ep_p2p_dispatch: bad arguments");
  if (int rc = b200moe_check_device()) return rc;
  PeerRows pr{reinterpret_cast<const unsigned char* const*>(peers), buf_offset_bytes, dest_of_row, rows_cap};
  ep_p2p_dispatch_kernel<<<p2p_grid(send_rows_cap), kP2pThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), token_of_row, send_offsets, send_meta, dest_off, pr, E, E_local, d,
      dest_of_row);
  return check_launch("ep_p2p_dispatch");
}

extern "C" int b200moe_ep_p2p_combine(const void* const* peers, int64_t buf_offset_bytes, int rows_cap,
                                      const int32_t* dest_of_row, const int32_t* row_of_assignment,
                                      const float* topk_gate, int T, int k, int d, void* y, void* stream) {
  B200MOE_REQUIRE(peers && dest_of_row && k >= 1 && k <= 8 && d % 8 == 0, B200MOE_EINVAL, "ep_p2p_combine: bad arguments");
  if (T == 0) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  PeerRows pr{reinterpret_cast<const unsigned char* const*>(peers), buf_offset_bytes, dest_of_row, rows_cap};
  ep_p2p_combine_kernel<<<p2p_grid(T), kP2pThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      pr, row_of_assignment, topk_gate, T, k, d, static_cast<__nv_bfloat16*>(y));
  return check_launch("ep_p2p_combine");
}

extern "C" int b200moe_ep_p2p_combine_bwd(const void* dy, const void* const* peers, int64_t yo_offset_bytes,
                                          int64_t dyp_offset_bytes, int rows_cap, const int32_t* dest_of_row,
                                          const int32_t* token_of_row, const float* gate_of_row,
                                          const int32_t* send_meta, int send_rows_cap, int d, float* dgate_row,
                                          void* stream) {
  B200MOE_REQUIRE(dy && peers && dest_of_row && dgate_row && d % 8 == 0, B200MOE_EINVAL, "ep_p2p_combine_bwd: bad arguments");
  if (int rc = b200moe_check_device()) return rc;
  const unsigned char* const* pp = reinterpret_cast<const unsigned char* const*>(peers);
  PeerRows yo{pp, yo_offset_bytes, dest_of_row, rows_cap}, dyp{pp, dyp_offset_bytes, dest_of_row, rows_cap};
  ep_p2p_combine_bwd_kernel<<<p2p_grid(send_rows_cap), kP2pThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), yo, dyp, token_of_row, gate_of_row, send_meta, d, dgate_row);
  return check_launch("ep_p2p_combine_bwd");
}

extern "C" int b200moe_ep_p2p_pull_rows(const void* const* peers, int64_t buf_offset_bytes, int rows_cap,
                                        const int32_t* dest_of_row, const int32_t* send_meta, int send_rows_cap, int d,
                                        void* out, void* stream) {
  B200MOE_REQUIRE(peers && dest_of_row && out && d % 8 == 0, B200MOE_EINVAL, "ep_p2p_pull_rows: bad arguments");
  if (int rc = b200moe_check_device()) return rc;
  PeerRows pr{reinterpret_cast<const unsigned char* const*>(peers), buf_offset_bytes, dest_of_row, rows_cap};
  ep_p2p_pull_rows_kernel<<<p2p_grid(send_rows_cap), kP2pThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      pr, send_meta, d, static_cast<__nv_bfloat16*>(out));
  return check_launch("ep_p2p_pull_rows");
}
