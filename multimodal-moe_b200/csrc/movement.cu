// HBM-bound row movement: dispatch (permute), combine, their backward passes, casts, column sums.
// One warp moves one row with 128-bit loads/stores (a row of d bf16 is d/8 16-byte chunks, lanes stride them),
// so every request is a run of full 32-byte sectors and consecutive lanes touch consecutive addresses.
#include <stdlib.h>

#include "common.cuh"
#include "postnorm.cuh"

namespace b200moe {

constexpr int kMoveWarps = 8;
constexpr int kMoveThreads = kMoveWarps * 32;

// What the dispatch kernel prepares for the down-projection's fused combine epilogue (gemm.cu, EPI_DOWN_COMBINE), so that
// no memset launch sits on the critical path: the fp32 accumulator and the grid rendezvous counter start at zero.
struct CombineInit {
  float* y_acc;        // [T,d] fp32 or null
  int32_t* grid_sync;  // 4 ints (16 bytes) or null
  int T;
};

// xp[r] = x[token_of_row[r]] or 0
__global__ void __launch_bounds__(kMoveThreads)
permute_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ token_of_row,
               const int32_t* __restrict__ meta, int d, __nv_bfloat16* __restrict__ xp, const CombineInit ci) {
  pdl_sync();
  const int R = meta[0];
  const int nchunk = d >> 3;
  const int lane = threadIdx.x & 31;
  for (int r = blockIdx.x * kMoveWarps + (threadIdx.x >> 5); r < R; r += gridDim.x * kMoveWarps) {
    const int t = token_of_row[r];
    __nv_bfloat16* dst = xp + static_cast<size_t>(r) * d;
    if (t >= 0) {
      const __nv_bfloat16* src = x + static_cast<size_t>(t) * d;
      for (int c = lane; c < nchunk; c += 32) stg128(dst + c * 8, ldg128(src + c * 8));
    } else {
      for (int c = lane; c < nchunk; c += 32) stg128(dst + c * 8, make_uint4(0, 0, 0, 0));
    }
  }
  const size_t gid = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (ci.grid_sync && gid == 0) stg128(ci.grid_sync, make_uint4(0, 0, 0, 0));
  if (ci.y_acc) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    const size_t n_acc = static_cast<size_t>(ci.T) * d / 4;             // 16-byte pieces
    for (size_t i = gid; i < n_acc; i += stride) stg128(ci.y_acc + 4 * i, make_uint4(0, 0, 0, 0));
  }
}

// out[src_of_row[r]] = in[r]   (EP return path)
__global__ void __launch_bounds__(kMoveThreads)
scatter_rows_kernel(const __nv_bfloat16* __restrict__ in, const int32_t* __restrict__ src_of_row,
                    const int32_t* __restrict__ meta, int d, __nv_bfloat16* __restrict__ out) {
  const int R = meta[0];
  const int nchunk = d >> 3;
  const int lane = threadIdx.x & 31;
  for (int r = blockIdx.x * kMoveWarps + (threadIdx.x >> 5); r < R; r += gridDim.x * kMoveWarps) {
    const int t = src_of_row[r];
    if (t < 0) continue;
    const __nv_bfloat16* src = in + static_cast<size_t>(r) * d;
    __nv_bfloat16* dst = out + static_cast<size_t>(t) * d;
    for (int c = lane; c < nchunk; c += 32) stg128(dst + c * 8, ldg128(src + c * 8));
  }
}

// y[t] = sum_s gate[t,s] * yo[row[t,s]]   (slot order, fp32)
__global__ void __launch_bounds__(kMoveThreads)
combine_kernel(const __nv_bfloat16* __restrict__ yo, const int32_t* __restrict__ row_of_assignment,
               const float* __restrict__ topk_gate, int T, int k, int d, __nv_bfloat16* __restrict__ y) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  for (int t = blockIdx.x * kMoveWarps + (threadIdx.x >> 5); t < T; t += gridDim.x * kMoveWarps)
    combine_token(yo, row_of_assignment, topk_gate, false, t, k, d, lane, y);
}

// dyp[r] = gate_of_row[r] * dy[token_of_row[r]];  dgate_row[r] = <dy[token], yo[r]>
__global__ void __launch_bounds__(kMoveThreads)
combine_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ yo,
                   const int32_t* __restrict__ token_of_row, const float* __restrict__ gate_of_row,
                   const int32_t* __restrict__ meta, int d, __nv_bfloat16* __restrict__ dyp,
                   float* __restrict__ dgate_row) {
  pdl_sync();
  const int R = meta[0];
  const int nchunk = d >> 3;
  const int lane = threadIdx.x & 31;
  for (int r = blockIdx.x * kMoveWarps + (threadIdx.x >> 5); r < R; r += gridDim.x * kMoveWarps) {
    const int t = token_of_row[r];
    __nv_bfloat16* dst = dyp + static_cast<size_t>(r) * d;
    if (t < 0) {
      for (int c = lane; c < nchunk; c += 32) stg128(dst + c * 8, make_uint4(0, 0, 0, 0));
      if (lane == 0 && dgate_row) dgate_row[r] = 0.f;
      continue;
    }
    const float g = gate_of_row[r];
    float dot = 0.f;
    for (int c = lane; c < nchunk; c += 32) {
      float a[8], b[8], o[8];
      unpack8(ldg128(dy + static_cast<size_t>(t) * d + c * 8), a);
      if (yo) unpack8(ldg128(yo + static_cast<size_t>(r) * d + c * 8), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        o[j] = g * a[j];
        if (yo) dot = fmaf(a[j], b[j], dot);
      }
      stg128(dst + c * 8, pack8(o));
    }
    if (dgate_row) {
      dot = warp_sum(dot);
      if (lane == 0) dgate_row[r] = dot;
    }
  }
}

// dx[t] = sum_s dxp[row[t,s]] + dlogits[t] @ wg      (backward of permute fused with the router's input gradient)
// Block = 64 tokens x one 256-column slab.  The slab of wg [E][256] fp32 and the block's dlogits [64][E] sit in
// shared memory; a warp owns 8 tokens, a lane two 4-column pieces (c, c+128: conflict-free 16-byte smem reads,
// 256 contiguous bytes per warp access in global memory), so every wg value read from shared memory feeds 8 FMAs
// per token x 8 tokens.  The k gathered rows of dxp are added in slot order in fp32.
// TT = tokens per warp (8 for large batches; 2 for small ones so that the grid still covers the SMs).
template <int TT, bool P2P>
__global__ void __launch_bounds__(256)
unpermute_bwd_kernel(const __nv_bfloat16* __restrict__ dxp, const int32_t* __restrict__ row_of_assignment,
                     const float* __restrict__ dlogits, const float* __restrict__ wg, int T, int k, int d, int E,
                     __nv_bfloat16* __restrict__ dx, const PeerRows pr) {
  constexpr int kTok = 8 * TT;         // tokens per block
  pdl_sync();
  extern __shared__ __align__(16) float unp_s[];
  float* wg_s = unp_s;                 // [E][256]
  float* dl_s = unp_s + E * 256;       // [kTok][E]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t0 = blockIdx.x * kTok;
  const int slab0 = blockIdx.y * 256;
  const int c0 = slab0 + lane * 4, c1 = c0 + 128;
  const bool has1 = c1 < d;            // d is a multiple of 128: the second piece is missing only when d == 128
  float acc[TT][8];
#pragma unroll
  for (int tt = 0; tt < TT; ++tt)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[tt][j] = 0.f;
  // gathers first (independent of the router part): all loads of up to 4 tokens in flight at once
  if (P2P || dxp) {
#pragma unroll
    for (int h = 0; h < TT; h += 4) {
      constexpr int kG = TT < 4 ? TT : 4;
      for (int s = 0; s < k; ++s) {
        uint2 va[kG], vb[kG];
#pragma unroll
        for (int q = 0; q < kG; ++q) {
          const int t = t0 + warp * TT + h + q;
          int r = t < T ? row_of_assignment[static_cast<size_t>(t) * k + s] : -1;
          const __nv_bfloat16* src;
          if (P2P) {                      // expert parallelism: the row lives on the expert's rank (ep_p2p.cu)
            r = r >= 0 ? pr.dest_of_row[r] : -1;
            src = r >= 0 ? peer_row(pr, r, d) : nullptr;
          } else {
            src = dxp + static_cast<size_t>(r < 0 ? 0 : r) * d;
          }
          va[q] = r >= 0 ? __ldg(reinterpret_cast<const uint2*>(src + c0)) : make_uint2(0, 0);
          vb[q] = (r >= 0 && has1) ? __ldg(reinterpret_cast<const uint2*>(src + c1)) : make_uint2(0, 0);
        }
#pragma unroll
        for (int q = 0; q < kG; ++q) {
          float2 f;
          f = bf16x2_to_float2(va[q].x); acc[h + q][0] += f.x; acc[h + q][1] += f.y;
          f = bf16x2_to_float2(va[q].y); acc[h + q][2] += f.x; acc[h + q][3] += f.y;
          f = bf16x2_to_float2(vb[q].x); acc[h + q][4] += f.x; acc[h + q][5] += f.y;
          f = bf16x2_to_float2(vb[q].y); acc[h + q][6] += f.x; acc[h + q][7] += f.y;
        }
      }
    }
  }
  if (dlogits) {
    for (int i = threadIdx.x; i < E * 64; i += 256) {
      const int e = i >> 6, c4 = (i & 63) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (slab0 + c4 < d) v = __ldg(reinterpret_cast<const float4*>(wg + static_cast<size_t>(e) * d + slab0 + c4));
      *reinterpret_cast<float4*>(wg_s + e * 256 + c4) = v;
    }
    const long long lim = static_cast<long long>(T) * E, base = static_cast<long long>(t0) * E;
    for (int i = threadIdx.x; i < kTok * E; i += 256) dl_s[i] = (base + i < lim) ? dlogits[base + i] : 0.f;
    __syncthreads();
    const float* dlw = dl_s + warp * TT * E;
#pragma unroll 2
    for (int e = 0; e < E; ++e) {
      const float4 wa = *reinterpret_cast<const float4*>(wg_s + e * 256 + lane * 4);
      const float4 wb = *reinterpret_cast<const float4*>(wg_s + e * 256 + 128 + lane * 4);
#pragma unroll
      for (int tt = 0; tt < TT; ++tt) {
        const float g = dlw[tt * E + e];
        acc[tt][0] = fmaf(g, wa.x, acc[tt][0]); acc[tt][1] = fmaf(g, wa.y, acc[tt][1]);
        acc[tt][2] = fmaf(g, wa.z, acc[tt][2]); acc[tt][3] = fmaf(g, wa.w, acc[tt][3]);
        acc[tt][4] = fmaf(g, wb.x, acc[tt][4]); acc[tt][5] = fmaf(g, wb.y, acc[tt][5]);
        acc[tt][6] = fmaf(g, wb.z, acc[tt][6]); acc[tt][7] = fmaf(g, wb.w, acc[tt][7]);
      }
    }
  }
#pragma unroll
  for (int tt = 0; tt < TT; ++tt) {
    const int t = t0 + warp * TT + tt;
    if (t < T) {
      __nv_bfloat16* dst = dx + static_cast<size_t>(t) * d;
      *reinterpret_cast<uint2*>(dst + c0) =
          make_uint2(float2_to_bf16x2(acc[tt][0], acc[tt][1]), float2_to_bf16x2(acc[tt][2], acc[tt][3]));
      if (has1)
        *reinterpret_cast<uint2*>(dst + c1) =
            make_uint2(float2_to_bf16x2(acc[tt][4], acc[tt][5]), float2_to_bf16x2(acc[tt][6], acc[tt][7]));
    }
  }
}

__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n8, int64_t n) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(src) + 2 * i);
    const float4 b = __ldg(reinterpret_cast<const float4*>(src) + 2 * i + 1);
    const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    stg128(dst + 8 * i, pack8(v));
  }
  if (blockIdx.x == 0)
    for (int64_t i = n8 * 8 + threadIdx.x; i < n; i += blockDim.x) dst[i] = __float2bfloat16_rn(src[i]);
}

// db[e, c] = sum over the rows of expert e of a[r, c] for up to two activation tensors in ONE launch, fixed order
// (deterministic, no atomics, no scratch).  Block = (64-column slab of one of the tensors, expert), 1024 threads:
// thread (chunk, phase) owns 8 columns and every 128th row of the expert's segment, four 16-byte loads in flight
// (64 KB per block), then the 128 phases are summed in a fixed two-level order through shared memory.
struct ColsumJob {
  const __nv_bfloat16* a;
  float* db;
  int ncols;
};
// Block = (64-column slab, expert), 256 threads: thread (phase, chunk) sums the rows phase, phase + 32, ... of its 8
// columns (four 16-byte loads in flight), the four phases of a warp are folded by shuffles, the eight warps through
// 512 bytes of shared memory in four rounds of 16 columns -- every sum in a fixed order.  The small footprint is the
// point: 256 threads x 40 registers and half a KB of shared memory fit on an SM BESIDE a persistent GEMM CTA (576
// threads x 96 registers, ~226 KB), so this bandwidth-bound kernel runs under the tensor-bound dgrad-up / weight-gradient
// GEMMs.  The round-1 version (1024 threads, 35 KB of shared memory) could not: it ran between them, alone on the GPU
// (round-2 timelines: a ~110 us hole in front of the weight gradients at C/65533 and at D).
constexpr int kColsumPhases = 32;

__global__ void __launch_bounds__(8 * kColsumPhases)
colsum_kernel(const ColsumJob j0, const ColsumJob j1, const int32_t* __restrict__ expert_offsets, int accumulate) {
  __shared__ float part[8][16];
  pdl_sync();
  const int slabs0 = j0.ncols >> 6;
  const bool second = static_cast<int>(blockIdx.x) >= slabs0;
  const ColsumJob j = second ? j1 : j0;
  const int slab = second ? blockIdx.x - slabs0 : blockIdx.x;
  const int e = blockIdx.y;
  const int r0 = expert_offsets[e], r1 = expert_offsets[e + 1];
  const int chunk = threadIdx.x & 7, phase = threadIdx.x >> 3;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int col0 = slab * 64 + chunk * 8;
  const __nv_bfloat16* base = j.a + col0;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int r = r0 + phase; r < r1; r += 4 * kColsumPhases) {
    uint4 raw[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int rr = r + kColsumPhases * u;
      raw[u] = rr < r1 ? ldg128(base + static_cast<size_t>(rr) * j.ncols) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float v[8];
      unpack8(raw[u], v);
#pragma unroll
      for (int q = 0; q < 8; ++q) acc[q] += v[q];
    }
  }
  // the four phases of a warp (lanes l, l+8, l+16, l+24 share a chunk): lanes 0..7 end up with the warp's sums
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], 8);
    acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], 16);
  }
  // the eight warps, 16 columns (two chunks) per round, in warp order
#pragma unroll 1
  for (int round = 0; round < 4; ++round) {
    if (lane < 8 && (chunk >> 1) == round) {
#pragma unroll
      for (int q = 0; q < 8; ++q) part[warp][(chunk & 1) * 8 + q] = acc[q];
    }
    __syncthreads();
    if (threadIdx.x < 16) {
      float sum = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) sum += part[w][threadIdx.x];
      float* o = j.db + static_cast<size_t>(e) * j.ncols + slab * 64 + round * 16 + threadIdx.x;
      *o = accumulate ? *o + sum : sum;
    }
    __syncthreads();
  }
}

// The round-1 shape of the same sum: 1024 threads per (slab, expert) and 35 KB of shared memory.  It cannot share an SM
// with a GEMM CTA, but it has four times the loads in flight per block: used when there are too few (slab, expert) blocks
// for 256 threads each to fill the memory system (8 experts x 20 slabs at 65600 tokens: 459 us with the small blocks,
// 80 us with these).
constexpr int kColsumWidePhases = 128;

__global__ void __launch_bounds__(8 * kColsumWidePhases)
colsum_wide_kernel(const ColsumJob j0, const ColsumJob j1, const int32_t* __restrict__ expert_offsets, int accumulate) {
  __shared__ float part[kColsumWidePhases][65];
  __shared__ float part2[8][64];
  pdl_sync();
  const int slabs0 = j0.ncols >> 6;
  const bool second = static_cast<int>(blockIdx.x) >= slabs0;
  const ColsumJob j = second ? j1 : j0;
  const int slab = second ? blockIdx.x - slabs0 : blockIdx.x;
  const int e = blockIdx.y;
  const int r0 = expert_offsets[e], r1 = expert_offsets[e + 1];
  const int chunk = threadIdx.x & 7, phase = threadIdx.x >> 3;
  const int col0 = slab * 64 + chunk * 8;
  const __nv_bfloat16* base = j.a + col0;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int r = r0 + phase; r < r1; r += 4 * kColsumWidePhases) {
    uint4 raw[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int rr = r + kColsumWidePhases * u;
      raw[u] = rr < r1 ? ldg128(base + static_cast<size_t>(rr) * j.ncols) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float v[8];
      unpack8(raw[u], v);
#pragma unroll
      for (int q = 0; q < 8; ++q) acc[q] += v[q];
    }
  }
#pragma unroll
  for (int q = 0; q < 8; ++q) part[phase][chunk * 8 + q] = acc[q];
  __syncthreads();
  if (threadIdx.x < 512) {
    const int g = threadIdx.x >> 6, c = threadIdx.x & 63;
    float sum = 0.f;
#pragma unroll
    for (int p = 0; p < 16; ++p) sum += part[g * 16 + p][c];
    part2[g][c] = sum;
  }
  __syncthreads();
  if (threadIdx.x < 64) {
    float sum = 0.f;
#pragma unroll
    for (int g = 0; g < 8; ++g) sum += part2[g][threadIdx.x];
    float* o = j.db + static_cast<size_t>(e) * j.ncols + slab * 64 + threadIdx.x;
    *o = accumulate ? *o + sum : sum;
  }
}

// Once per optimizer step, ONE launch: fp32 master weights of the experts -> bf16 operands, and the gate matrix
// -> its bf16 hi/lo split for the tcgen05 router (rows 0..E-1 hi, rows EP..EP+E-1 lo = bf16(w - hi), rest zero).
__global__ void __launch_bounds__(256)
prepare_weights_kernel(const float* __restrict__ w1, __nv_bfloat16* __restrict__ w1b, int64_t n1_8,
                       const float* __restrict__ w2, __nv_bfloat16* __restrict__ w2b, int64_t n2_8,
                       const float* __restrict__ wg, __nv_bfloat16* __restrict__ wsplit, int E, int EP, int d) {
  const int64_t ns_8 = wg ? (static_cast<int64_t>(2) * EP * d) >> 3 : 0;
  const int64_t total = n1_8 + n2_8 + ns_8;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int64_t nw_8 = n1_8 + n2_8;
  // expert slabs: two 32-byte reads per thread in flight (the grid leaves half of every SM's thread slots to the
  // kernels this cast runs beside: with 2048 resident threads per SM the router waited for the cast's last wave --
  // 204 us at 32 local experts of workload D, round-2 timeline)
  int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; i + stride < nw_8; i += 2 * stride) {
    const int64_t i1 = i + stride;
    const bool f0 = i < n1_8, f1 = i1 < n1_8;
    const float4* s0 = reinterpret_cast<const float4*>(f0 ? w1 : w2) + 2 * (f0 ? i : i - n1_8);
    const float4* s1 = reinterpret_cast<const float4*>(f1 ? w1 : w2) + 2 * (f1 ? i1 : i1 - n1_8);
    const float4 a0 = __ldg(s0), b0 = __ldg(s0 + 1), a1 = __ldg(s1), b1 = __ldg(s1 + 1);
    const float v0[8] = {a0.x, a0.y, a0.z, a0.w, b0.x, b0.y, b0.z, b0.w};
    const float v1[8] = {a1.x, a1.y, a1.z, a1.w, b1.x, b1.y, b1.z, b1.w};
    stg128((f0 ? w1b : w2b) + 8 * (f0 ? i : i - n1_8), pack8(v0));
    stg128((f1 ? w1b : w2b) + 8 * (f1 ? i1 : i1 - n1_8), pack8(v1));
  }
  for (; i < total; i += stride) {
    if (i < nw_8) {
      const bool first = i < n1_8;
      const float* src = first ? w1 : w2;
      __nv_bfloat16* dst = first ? w1b : w2b;
      const int64_t c = first ? i : i - n1_8;
      const float4 a = __ldg(reinterpret_cast<const float4*>(src) + 2 * c);
      const float4 b = __ldg(reinterpret_cast<const float4*>(src) + 2 * c + 1);
      const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
      stg128(dst + 8 * c, pack8(v));
    } else {
      const int64_t c = i - nw_8;
      const int64_t el = c * 8;
      const int row = static_cast<int>(el / d), col = static_cast<int>(el - static_cast<int64_t>(row) * d);
      const int e = row < EP ? row : row - EP;
      float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (e < E) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(wg + static_cast<size_t>(e) * d + col));
        const float4 b = __ldg(reinterpret_cast<const float4*>(wg + static_cast<size_t>(e) * d + col) + 1);
        const float w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int q = 0; q < 8; ++q)
          v[q] = row < EP ? w[q] : w[q] - __bfloat162float(__float2bfloat16_rn(w[q]));
      }
      stg128(wsplit + el, pack8(v));
    }
  }
}

// ---- fused post-norm (hybrid_encoder.py:158-160): out = LayerNorm(residual + combine(yo)) ---------------------------
// One warp per token; a lane owns the 16-byte chunks lane, lane+32, ... of the row (d <= 1024: at most four), so the
// gate-weighted sum, the residual add, both LayerNorm moments and the affine output never leave registers and the
// token's [T,d] intermediate (y, y + residual) is never written.  mean / rstd are kept for the backward pass.
template <int NI>
__global__ void __launch_bounds__(kMoveThreads)
combine_ln_fwd_kernel(const __nv_bfloat16* __restrict__ yo, const int32_t* __restrict__ roa,
                      const float* __restrict__ gate, const float* __restrict__ y_acc,
                      const void* __restrict__ res, int res_f32,
                      const float* __restrict__ gamma, const float* __restrict__ beta, float eps, int T, int k, int d,
                      void* __restrict__ out, int out_f32, float* __restrict__ mean_rstd) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  for (int t = blockIdx.x * kMoveWarps + (threadIdx.x >> 5); t < T; t += gridDim.x * kMoveWarps)
    ln_forward_token<NI>(yo, roa, gate, y_acc, false, res, res_f32 != 0, gamma, beta, eps, t, k, d, lane, out,
                         out_f32 != 0, mean_rstd);
}

// Backward: recomputes the pre-norm sum v and xhat from the same inputs, dv = rstd (dxhat - mean(dxhat) - xhat
// mean(dxhat xhat)) with dxhat = dout gamma; dv is both the residual's gradient and the MoE output's gradient.
// dgamma / dbeta: per-lane partial sums over the warp's tokens, the block's warps in fixed order, one row of scratch
// per block; a second launch sums the blocks in order (deterministic).
template <int NI>
__global__ void __launch_bounds__(kMoveThreads)
combine_ln_bwd_kernel(const void* __restrict__ dout, int dout_f32, const __nv_bfloat16* __restrict__ yo,
                      const int32_t* __restrict__ roa, const float* __restrict__ gate,
                      const float* __restrict__ y_acc,
                      const void* __restrict__ res, int res_f32, const float* __restrict__ gamma,
                      const float* __restrict__ mean_rstd, int T, int k, int d, __nv_bfloat16* __restrict__ dv,
                      float* __restrict__ dres_f32, float* __restrict__ scratch) {
  extern __shared__ __align__(16) float lnb_s[];       // [8 warps][2][d]
  pdl_sync();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nchunk = d >> 3;
  const float invd = 1.0f / static_cast<float>(d);
  float dg[NI][8], db[NI][8];
#pragma unroll
  for (int i = 0; i < NI; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) { dg[i][j] = 0.f; db[i][j] = 0.f; }
  for (int t = blockIdx.x * kMoveWarps + warp; t < T; t += gridDim.x * kMoveWarps) {
    float v[NI][8], dy[NI][8];
    ln_gather_sum<NI>(yo, roa, gate, y_acc, false, res, res_f32 != 0, t, k, d, lane, v);
    const float mean = mean_rstd[2 * static_cast<size_t>(t)], rstd = mean_rstd[2 * static_cast<size_t>(t) + 1];
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
      const int c = lane + 32 * i;
      if (c < nchunk) {
        float a[8];
        load8_any(dout, dout_f32 != 0, static_cast<size_t>(t) * d + c * 8, a);
        const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c * 8)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c * 8) + 1);
        const float gm[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = (v[i][j] - mean) * rstd;
          v[i][j] = xh;
          dg[i][j] = fmaf(a[j], xh, dg[i][j]);
          db[i][j] += a[j];
          dy[i][j] = a[j] * gm[j];
          m1 += dy[i][j];
          m2 = fmaf(dy[i][j], xh, m2);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) dy[i][j] = 0.f;
      }
    }
    m1 = warp_sum(m1) * invd;
    m2 = warp_sum(m2) * invd;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
      const int c = lane + 32 * i;
      if (c < nchunk) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (dy[i][j] - m1 - v[i][j] * m2);
        stg128(dv + static_cast<size_t>(t) * d + c * 8, pack8(o));
        if (dres_f32) store8_any(dres_f32, true, static_cast<size_t>(t) * d + c * 8, o);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < NI; ++i) {
    const int c = lane + 32 * i;
    if (c < nchunk) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        lnb_s[(warp * 2 + 0) * d + c * 8 + j] = dg[i][j];
        lnb_s[(warp * 2 + 1) * d + c * 8 + j] = db[i][j];
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * d; i += blockDim.x) {
    float sum = 0.f;
#pragma unroll
    for (int w = 0; w < kMoveWarps; ++w) sum += lnb_s[w * 2 * d + i];
    scratch[static_cast<size_t>(blockIdx.x) * 2 * d + i] = sum;
  }
}

__global__ void __launch_bounds__(256)
ln_param_reduce_kernel(const float* __restrict__ scratch, int nblk, int d, float* __restrict__ dgamma,
                       float* __restrict__ dbeta) {
  pdl_sync();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2 * d) return;
  float s = 0.f;
  for (int b = 0; b < nblk; ++b) s += scratch[static_cast<size_t>(b) * 2 * d + i];
  if (i < d) dgamma[i] = s;
  else dbeta[i - d] = s;
}

static inline int ln_bwd_blocks(int T) {
  int b = (T + kMoveWarps - 1) / kMoveWarps;
  if (b > num_sms() * 2) b = num_sms() * 2;
  return b < 1 ? 1 : b;
}

static inline int move_grid(long long rows) {
  long long b = (rows + kMoveWarps - 1) / kMoveWarps;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace b200moe

using namespace b200moe;
#define BF(p) static_cast<const __nv_bfloat16*>(p)
#define BFW(p) static_cast<__nv_bfloat16*>(p)
#define ST(s) static_cast<cudaStream_t>(s)

extern "C" int b200moe_permute(const void* x, const int32_t* token_of_row, const int32_t* meta, int rows_cap, int d,
                               void* xp, void* stream) {
  return b200moe_permute_init(x, token_of_row, meta, rows_cap, d, xp, nullptr, nullptr, 0, stream);
}

extern "C" int b200moe_permute_init(const void* x, const int32_t* token_of_row, const int32_t* meta, int rows_cap,
                                    int d, void* xp, float* y_acc, int32_t* grid_sync, int T, void* stream) {
  B200MOE_REQUIRE(d > 0 && d % 8 == 0, B200MOE_EINVAL, "permute: d=%d must be a multiple of 8", d);
  B200MOE_REQUIRE(y_acc == nullptr || (T > 0 && d % 4 == 0), B200MOE_EINVAL, "permute_init: bad T=%d / d=%d", T, d);
  B200MOE_REQUIRE(grid_sync == nullptr || (reinterpret_cast<uintptr_t>(grid_sync) & 15) == 0, B200MOE_EINVAL,
                  "permute_init: grid_sync needs 16-byte alignment");
  if (int rc = b200moe_check_device()) return rc;
  const CombineInit ci{y_acc, grid_sync, T};
  launch_k(permute_kernel, move_grid(rows_cap), kMoveThreads, 0, ST(stream), BF(x), token_of_row, meta, d, BFW(xp), ci);
  return check_launch("permute");
}

extern "C" int b200moe_scatter_rows(const void* in, const int32_t* src_of_row, const int32_t* meta, int rows_cap,
                                    int d, void* out, void* stream) {
  B200MOE_REQUIRE(d > 0 && d % 8 == 0, B200MOE_EINVAL, "scatter_rows: d=%d must be a multiple of 8", d);
  if (int rc = b200moe_check_device()) return rc;
  scatter_rows_kernel<<<move_grid(rows_cap), kMoveThreads, 0, ST(stream)>>>(BF(in), src_of_row, meta, d, BFW(out));
  return check_launch("scatter_rows");
}

extern "C" int b200moe_combine(const void* yo, const int32_t* row_of_assignment, const float* topk_gate, int T,
                               int k, int d, void* y, void* stream) {
  B200MOE_REQUIRE(d > 0 && d % 8 == 0, B200MOE_EINVAL, "combine: d=%d must be a multiple of 8", d);
  if (T == 0) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  launch_k(combine_kernel, move_grid(T), kMoveThreads, 0, ST(stream), BF(yo), row_of_assignment, topk_gate, T, k, d, BFW(y));
  return check_launch("combine");
}

extern "C" int b200moe_combine_bwd(const void* dy, const void* yo, const int32_t* token_of_row,
                                   const float* gate_of_row, const int32_t* meta, int rows_cap, int d, void* dyp,
                                   float* dgate_row, void* stream) {
  B200MOE_REQUIRE(d > 0 && d % 8 == 0, B200MOE_EINVAL, "combine_bwd: d=%d must be a multiple of 8", d);
  if (int rc = b200moe_check_device()) return rc;
  launch_k(combine_bwd_kernel, move_grid(rows_cap), kMoveThreads, 0, ST(stream), BF(dy), BF(yo), token_of_row,
           gate_of_row, meta, d, BFW(dyp), dgate_row);
  return check_launch("combine_bwd");
}

static int unpermute_bwd_impl(const void* dxp, const b200moe::PeerRows& pr, const int32_t* row_of_assignment,
                              const float* dlogits, const float* wg, int T, int k, int d, int E, void* dx,
                              void* stream) {
  using namespace b200moe;
  B200MOE_REQUIRE(d > 0 && d % 128 == 0, B200MOE_EINVAL, "unpermute_bwd: d=%d must be a multiple of 128", d);
  B200MOE_REQUIRE(E >= 1 && E <= 64, B200MOE_EINVAL, "unpermute_bwd: num_experts=%d outside [1,64]", E);
  B200MOE_REQUIRE(dlogits == nullptr || wg != nullptr, B200MOE_EINVAL, "unpermute_bwd: dlogits without wg");
  if (T == 0) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  // 8 tokens per warp (each shared-memory read of wg feeds 64 FMAs) only where the router term is the limit: many
  // experts and enough tokens that 64-token blocks still cover the SMs; below that the kernel is latency-bound and
  // small blocks win (tools/unp_bench.py: 76 vs 86 us at E=16 / 65533 tokens, 88 vs 141 us at E=64 / 16384)
  const bool big = T >= 64 * num_sms() && E > 16;
  static const int tt_env = [] { const char* e = getenv("B200MOE_UNP_TT"); return e ? atoi(e) : 0; }();   // experiments
  const int tt = (tt_env == 2 || tt_env == 4 || tt_env == 8) ? tt_env : (big ? 8 : 2);
  const int tok = 8 * tt;
  const size_t smem = dlogits ? sizeof(float) * (static_cast<size_t>(E) * 256 + tok * E) : 0;
  static DeviceOnce unp_once;
  const int dev = current_device();
  if (!unp_once.done(dev)) {
    const int mx = static_cast<int>(sizeof(float) * (64 * 256 + 64 * 64));
    cudaError_t e1 = cudaFuncSetAttribute(unpermute_bwd_kernel<8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx);
    cudaError_t e2 = cudaFuncSetAttribute(unpermute_bwd_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx);
    cudaError_t e3 = cudaFuncSetAttribute(unpermute_bwd_kernel<8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx);
    cudaError_t e4 = cudaFuncSetAttribute(unpermute_bwd_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx);
    cudaFuncSetAttribute(unpermute_bwd_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx);
    cudaFuncSetAttribute(unpermute_bwd_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx);
    B200MOE_REQUIRE(e1 == cudaSuccess && e2 == cudaSuccess && e3 == cudaSuccess && e4 == cudaSuccess, B200MOE_ECUDA,
                    "unpermute_bwd: cudaFuncSetAttribute failed");
    unp_once.set(dev);
  }
  dim3 grid((T + tok - 1) / tok, (d + 255) / 256);
  const bool p2p = pr.peers != nullptr;
#define UNP_LAUNCH(TT_, P_)                                                                                          \
  launch_k(unpermute_bwd_kernel<TT_, P_>, grid, 256, smem, ST(stream), BF(dxp), row_of_assignment, dlogits, wg, T, k, \
           d, E, BFW(dx), pr)
  if (tt == 8 && p2p) UNP_LAUNCH(8, true);
  else if (tt == 8) UNP_LAUNCH(8, false);
  else if (tt == 4 && p2p) UNP_LAUNCH(4, true);
  else if (tt == 4) UNP_LAUNCH(4, false);
  else if (p2p) UNP_LAUNCH(2, true);
  else UNP_LAUNCH(2, false);
#undef UNP_LAUNCH
  return check_launch("unpermute_bwd");
}

extern "C" int b200moe_unpermute_bwd(const void* dxp, const int32_t* row_of_assignment, const float* dlogits,
                                     const float* wg, int T, int k, int d, int E, void* dx, void* stream) {
  return unpermute_bwd_impl(dxp, b200moe::PeerRows{nullptr, 0, nullptr, 1}, row_of_assignment, dlogits, wg, T, k, d, E,
                            dx, stream);
}

extern "C" int b200moe_ep_p2p_unpermute_bwd(const void* const* peers, int64_t buf_offset_bytes, int rows_cap,
                                            const int32_t* dest_of_row, const int32_t* row_of_assignment,
                                            const float* dlogits, const float* wg, int T, int k, int d, int E,
                                            void* dx, void* stream) {
  B200MOE_REQUIRE(peers && dest_of_row && rows_cap > 0, B200MOE_EINVAL, "ep_p2p_unpermute_bwd: bad arguments");
  b200moe::PeerRows pr{reinterpret_cast<const unsigned char* const*>(peers), buf_offset_bytes, dest_of_row, rows_cap};
  return unpermute_bwd_impl(nullptr, pr, row_of_assignment, dlogits, wg, T, k, d, E, dx, stream);
}

extern "C" int b200moe_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream) {
  B200MOE_REQUIRE(n >= 0, B200MOE_EINVAL, "cast: negative length");
  if (n == 0) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  const int64_t n8 = n / 8;
  int64_t blocks = (n8 + 255) / 256;
  if (blocks > num_sms() * 8) blocks = num_sms() * 8;
  if (blocks < 1) blocks = 1;
  cast_f32_bf16_kernel<<<static_cast<int>(blocks), 256, 0, ST(stream)>>>(src, BFW(dst), n8, n);
  return check_launch("cast_f32_bf16");
}

extern "C" int b200moe_colsum(const void* a0, int ncols0, float* db0, const void* a1, int ncols1, float* db1,
                              const int32_t* expert_offsets, int E, int accumulate, void* stream) {
  B200MOE_REQUIRE(a0 != nullptr && db0 != nullptr && ncols0 > 0 && ncols0 % 64 == 0, B200MOE_EINVAL,
                  "colsum: ncols=%d must be a positive multiple of 64", ncols0);
  B200MOE_REQUIRE(a1 == nullptr || (db1 != nullptr && ncols1 > 0 && ncols1 % 64 == 0), B200MOE_EINVAL,
                  "colsum: second tensor: ncols=%d must be a positive multiple of 64", ncols1);
  B200MOE_REQUIRE(E >= 1 && E <= 64, B200MOE_EINVAL, "colsum: num_experts=%d outside [1,64]", E);
  if (int rc = b200moe_check_device()) return rc;
  ColsumJob j0{BF(a0), db0, ncols0};
  ColsumJob j1{BF(a1), db1, a1 ? ncols1 : 0};
  dim3 grid((ncols0 >> 6) + (a1 ? (ncols1 >> 6) : 0), E);
  // small co-resident blocks once there are enough of them (blocks x 256 threads >= one full wave of 1024 per SM)
  if (static_cast<long long>(grid.x) * grid.y >= 4LL * num_sms())
    launch_k(colsum_kernel, grid, 8 * kColsumPhases, 0, ST(stream), j0, j1, expert_offsets, accumulate);
  else
    launch_k(colsum_wide_kernel, grid, 8 * kColsumWidePhases, 0, ST(stream), j0, j1, expert_offsets, accumulate);
  return check_launch("colsum");
}

extern "C" int b200moe_prepare_weights(const float* w1, void* w1_bf16, int64_t n1, const float* w2, void* w2_bf16,
                                       int64_t n2, const float* wg, void* wg_split, int d, int E, void* stream) {
  B200MOE_REQUIRE(n1 >= 0 && n2 >= 0 && n1 % 8 == 0 && n2 % 8 == 0, B200MOE_EINVAL,
                  "prepare_weights: element counts must be multiples of 8");
  B200MOE_REQUIRE(wg == nullptr || (wg_split != nullptr && d > 0 && d % 8 == 0 && E >= 1 && E <= 64), B200MOE_EINVAL,
                  "prepare_weights: bad gate shape d=%d E=%d", d, E);
  if (int rc = b200moe_check_device()) return rc;
  const int EP = E <= 16 ? 16 : (E <= 32 ? 32 : 64);
  const int64_t total = n1 / 8 + n2 / 8 + (wg ? (static_cast<int64_t>(2) * EP * d) / 8 : 0);
  if (total == 0) return B200MOE_OK;
  int64_t blocks = (total + 255) / 256;
  // Blocks per SM.  4 = half of the thread slots, so that the routing / dispatch kernels this cast runs beside find room
  // (expert parallelism, 32 local experts of workload D: 2.34 ms per step against 2.42 with 8).  A cast longer than
  // anything it can hide behind (>= 2.4 GB of traffic ~ 0.4 ms: all 64 experts of workload D on one GPU) takes every
  // slot instead: the up-projection waits for it either way, and it finishes in 0.56 ms instead of 0.73 (2.70 vs 2.86 ms
  // per step).  B200MOE_PREP_BLOCKS_PER_SM: measurement aid.
  static const int env_per_sm = [] { const char* e = getenv("B200MOE_PREP_BLOCKS_PER_SM"); const int v = e ? atoi(e) : 0; return v >= 1 && v <= 8 ? v : 0; }();
  const int per_sm = env_per_sm ? env_per_sm : ((n1 + n2) * 6 >= 2400000000ll ? 8 : 4);
  if (blocks > static_cast<int64_t>(num_sms()) * per_sm) blocks = static_cast<int64_t>(num_sms()) * per_sm;
  prepare_weights_kernel<<<static_cast<int>(blocks), 256, 0, ST(stream)>>>(
      w1, BFW(w1_bf16), n1 / 8, w2, BFW(w2_bf16), n2 / 8, wg, BFW(wg_split), E, EP, d);
  return check_launch("prepare_weights");
}

extern "C" int b200moe_combine_ln_fwd(const void* yo, const int32_t* row_of_assignment, const float* topk_gate,
                                      const float* y_acc, const void* residual, int residual_f32, const float* gamma,
                                      const float* beta, float eps, int T, int k, int d, void* out, int out_f32,
                                      float* mean_rstd, void* stream) {
  B200MOE_REQUIRE(d > 0 && d % 8 == 0 && d <= 1024, B200MOE_EINVAL, "combine_ln_fwd: d=%d must be a multiple of 8, <= 1024", d);
  B200MOE_REQUIRE((y_acc || (yo && row_of_assignment && topk_gate)) && residual && gamma && beta && out && mean_rstd,
                  B200MOE_EINVAL, "combine_ln_fwd: NULL argument");
  if (T == 0) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  const int ni = (d / 8 + 31) / 32;
#define LNF(NI)                                                                                                       \
  launch_k(combine_ln_fwd_kernel<NI>, move_grid(T), kMoveThreads, 0, ST(stream), BF(yo), row_of_assignment, topk_gate, \
           y_acc, residual, residual_f32, gamma, beta, eps, T, k, d, out, out_f32, mean_rstd)
  if (ni == 1) LNF(1);
  else if (ni == 2) LNF(2);
  else LNF(4);
#undef LNF
  return check_launch("combine_ln_fwd");
}

extern "C" int64_t b200moe_ln_scratch_floats(int T, int d) { return static_cast<int64_t>(ln_bwd_blocks(T)) * 2 * d; }

extern "C" int b200moe_combine_ln_bwd(const void* dout, int dout_f32, const void* yo, const int32_t* row_of_assignment,
                                      const float* topk_gate, const float* y_acc, const void* residual,
                                      int residual_f32, const float* gamma, const float* mean_rstd, int T, int k, int d,
                                      void* dv, float* dres_f32, float* scratch, float* dgamma, float* dbeta,
                                      void* stream) {
  B200MOE_REQUIRE(d > 0 && d % 8 == 0 && d <= 1024, B200MOE_EINVAL, "combine_ln_bwd: d=%d must be a multiple of 8, <= 1024", d);
  B200MOE_REQUIRE(dout && (y_acc || (yo && row_of_assignment && topk_gate)) && residual && gamma && mean_rstd && dv &&
                      scratch && dgamma && dbeta, B200MOE_EINVAL, "combine_ln_bwd: NULL argument");
  if (T == 0) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  const int ni = (d / 8 + 31) / 32, nblk = ln_bwd_blocks(T);
  const size_t smem = sizeof(float) * kMoveWarps * 2 * d;
  static DeviceOnce lnb_once;
  const int dev = current_device();
  if (!lnb_once.done(dev)) {
    const int mx = static_cast<int>(sizeof(float) * kMoveWarps * 2 * 1024);
    cudaError_t e1 = cudaFuncSetAttribute(combine_ln_bwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx);
    cudaError_t e2 = cudaFuncSetAttribute(combine_ln_bwd_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx);
    cudaError_t e3 = cudaFuncSetAttribute(combine_ln_bwd_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx);
    B200MOE_REQUIRE(e1 == cudaSuccess && e2 == cudaSuccess && e3 == cudaSuccess, B200MOE_ECUDA, "combine_ln_bwd: cudaFuncSetAttribute failed");
    lnb_once.set(dev);
  }
#define LNB(NI)                                                                                                     \
  launch_k(combine_ln_bwd_kernel<NI>, nblk, kMoveThreads, smem, ST(stream), dout, dout_f32, BF(yo), row_of_assignment, \
           topk_gate, y_acc, residual, residual_f32, gamma, mean_rstd, T, k, d, BFW(dv), dres_f32, scratch)
  if (ni == 1) LNB(1);
  else if (ni == 2) LNB(2);
  else LNB(4);
#undef LNB
  if (int rc = check_launch("combine_ln_bwd")) return rc;
  launch_k(ln_param_reduce_kernel, (2 * d + 255) / 256, 256, 0, ST(stream), static_cast<const float*>(scratch), nblk, d,
           dgamma, dbeta);
  return check_launch("ln_param_reduce");
}
