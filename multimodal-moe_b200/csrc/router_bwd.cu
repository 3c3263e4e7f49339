// Router gradient.
//   dL/dlogits[t,:] has three sources, all evaluated in fp32 from the saved logits:
//     gates : g_s = p_s / (S + eps), S = sum of the selected p  ->  dL/dp_{i_s} = (dg_s - sum_j dg_j g_j) / (S + eps)
//             (raw p when normalize_topk == 0); dropped slots have dg = 0 but stay in the normaliser
//     L_lb  : E * sum_e f_e P_e, f constant                      ->  dL/dp[t,e] += d_lb * E * f_e / T
//     L_z   : mean_t lse^2                                        ->  dL/dlogit[t,e] += d_z * 2 lse_t p[t,e] / T
//   followed by the softmax Jacobian  dlogit_e = p_e (dp_e - sum_j p_j dp_j).
// Then dWg = dlogits^T x (two-stage fixed-order reduction), dctx_bias = per-image segment sums.
#include <math.h>

#include "common.cuh"

namespace b200moe {

template <int EP>
__global__ void __launch_bounds__(128)
router_dlogits_kernel(const float* __restrict__ logits, const int32_t* __restrict__ topk_idx,
                      const int32_t* __restrict__ row_of_assignment, const float* __restrict__ dgate_row,
                      const int32_t* __restrict__ tokens_per_expert, const float* __restrict__ d_aux, int T, int E,
                      int k, int normalize, float aux_w, float z_w, float* __restrict__ dlogits) {
  __shared__ float f_s[64];
  if (threadIdx.x < E)
    f_s[threadIdx.x] = static_cast<float>(tokens_per_expert[threadIdx.x]) / static_cast<float>(static_cast<long long>(T) * k);
  __syncthreads();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const float da = d_aux ? *d_aux : 0.f;
  const float d_lb = da * aux_w, d_z = da * z_w;
  const float invT = 1.0f / static_cast<float>(T);

  float l[EP], p[EP], dp[EP];
  float mx = -INFINITY;
#pragma unroll
  for (int e = 0; e < EP; ++e) {
    l[e] = (e < E) ? logits[static_cast<size_t>(t) * E + e] : -INFINITY;
    mx = fmaxf(mx, l[e]);
  }
  float ssum = 0.f;
#pragma unroll
  for (int e = 0; e < EP; ++e) {
    p[e] = (e < E) ? expf(l[e] - mx) : 0.f;
    ssum += p[e];
  }
  const float inv = 1.0f / ssum;
  const float lse = mx + logf(ssum);
#pragma unroll
  for (int e = 0; e < EP; ++e) {
    p[e] *= inv;
    dp[e] = (e < E) ? d_lb * static_cast<float>(E) * f_s[e] * invT : 0.f;
  }
  // gate path
  if (dgate_row) {
    int idx[8];
    float dg[8], ps[8];
    float S = 0.f;
    for (int s = 0; s < k; ++s) {
      idx[s] = topk_idx[static_cast<size_t>(t) * k + s];
      const int r = row_of_assignment[static_cast<size_t>(t) * k + s];
      dg[s] = r >= 0 ? dgate_row[r] : 0.f;
      float pv = 0.f;
#pragma unroll
      for (int e = 0; e < EP; ++e) pv = (e == idx[s]) ? p[e] : pv;
      ps[s] = pv;
      S += pv;
    }
    float corr = 0.f;
    const float den = S + 1e-9f;
    if (normalize)
      for (int s = 0; s < k; ++s) corr += dg[s] * (ps[s] / den);
    for (int s = 0; s < k; ++s) {
      const float g = normalize ? (dg[s] - corr) / den : dg[s];
#pragma unroll
      for (int e = 0; e < EP; ++e) dp[e] += (e == idx[s]) ? g : 0.f;
    }
  }
  float dot = 0.f;
#pragma unroll
  for (int e = 0; e < EP; ++e) dot = fmaf(p[e], dp[e], dot);
  const float zc = d_z * 2.0f * lse * invT;
#pragma unroll
  for (int e = 0; e < EP; ++e)
    if (e < E) dlogits[static_cast<size_t>(t) * E + e] = p[e] * (dp[e] - dot) + zc * p[e];
}

// stage 1: block (token chunk c, column slab j) -> scratch[c][e][col] ; thread owns one column and E accumulators
constexpr int kDwgTok = 64;   // minimum tokens per chunk
template <int EP>
__global__ void __launch_bounds__(256)
router_dwg_partial_kernel(const float* __restrict__ dlogits, const __nv_bfloat16* __restrict__ x, int T, int d, int E,
                          int tok_per_chunk, float* __restrict__ scratch) {
  __shared__ float dl_s[64][EP];
  const int col = blockIdx.y * 256 + threadIdx.x;
  const int t_begin = blockIdx.x * tok_per_chunk;
  const int t_end = min(T, t_begin + tok_per_chunk);
  float acc[EP];
#pragma unroll
  for (int e = 0; e < EP; ++e) acc[e] = 0.f;
  for (int tb = t_begin; tb < t_end; tb += 64) {
    const int nt = min(64, t_end - tb);
    __syncthreads();
    for (int i = threadIdx.x; i < nt * E; i += blockDim.x) dl_s[i / E][i % E] = dlogits[static_cast<size_t>(tb) * E + i];
    __syncthreads();
    if (col < d) {
      for (int i = 0; i < nt; ++i) {
        const float xv = __bfloat162float(x[static_cast<size_t>(tb + i) * d + col]);
#pragma unroll
        for (int e = 0; e < EP; ++e)
          if (e < E) acc[e] = fmaf(dl_s[i][e], xv, acc[e]);
      }
    }
  }
  if (col < d) {
#pragma unroll
    for (int e = 0; e < EP; ++e)
      if (e < E) scratch[(static_cast<size_t>(blockIdx.x) * E + e) * d + col] = acc[e];
  }
}

__global__ void __launch_bounds__(256)
router_dwg_reduce_kernel(const float* __restrict__ scratch, int nchunks, int n, float* __restrict__ dwg) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float s = 0.f;
  for (int c = 0; c < nchunks; ++c) s += scratch[static_cast<size_t>(c) * n + i];
  dwg[i] = s;
}

// dctx[b, e] = sum over the tokens of image b.  One block per image, warp w handles experts w, w+8, ...
__global__ void __launch_bounds__(256)
router_dctx_kernel(const float* __restrict__ dlogits, int T, int E, int tpi, float* __restrict__ dctx) {
  const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t0 = b * tpi, t1 = min(T, t0 + tpi);
  for (int e = warp; e < E; e += 8) {
    float s = 0.f;
    for (int t = t0 + lane; t < t1; t += 32) s += dlogits[static_cast<size_t>(t) * E + e];
    s = warp_sum(s);
    if (lane == 0) dctx[static_cast<size_t>(b) * E + e] = s;
  }
}

static int dwg_chunks(int T) {
  int chunks = (T + kDwgTok - 1) / kDwgTok;
  const int cap = kNumSMs * 4;
  if (chunks > cap) chunks = cap;
  return chunks < 1 ? 1 : chunks;
}

}  // namespace b200moe

using namespace b200moe;

extern "C" int b200moe_router_bwd_dlogits(const float* logits, const int32_t* topk_idx,
                                          const int32_t* row_of_assignment, const float* dgate_row,
                                          const int32_t* tokens_per_expert, const float* d_aux, int T, int E, int k,
                                          int normalize_topk, float aux_loss_weight, float z_loss_weight,
                                          float* dlogits, void* stream) {
  B200MOE_REQUIRE(E >= 1 && E <= 64 && k >= 1 && k <= 8 && k <= E, B200MOE_EINVAL, "router_bwd: bad E=%d / k=%d", E, k);
  if (T == 0) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = (T + 127) / 128;
#define CASE(EP)                                                                                                  \
  router_dlogits_kernel<EP><<<grid, 128, 0, st>>>(logits, topk_idx, row_of_assignment, dgate_row, tokens_per_expert, \
                                                  d_aux, T, E, k, normalize_topk, aux_loss_weight, z_loss_weight, dlogits)
  if (E <= 8) CASE(8);
  else if (E <= 16) CASE(16);
  else if (E <= 32) CASE(32);
  else CASE(64);
#undef CASE
  return check_launch("router_dlogits");
}

extern "C" int64_t b200moe_dwg_scratch_floats(int T, int d, int E) {
  return static_cast<int64_t>(dwg_chunks(T)) * E * d;
}

extern "C" int b200moe_router_bwd_dwg(const float* dlogits, const void* x, int T, int d, int E, float* scratch,
                                      float* dwg, void* stream) {
  B200MOE_REQUIRE(E >= 1 && E <= 64 && d > 0, B200MOE_EINVAL, "router_bwd_dwg: bad E=%d / d=%d", E, d);
  if (int rc = b200moe_check_device()) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int chunks = dwg_chunks(T);
  const int tpc = (T + chunks - 1) / chunks;
  dim3 grid(chunks, (d + 255) / 256);
  auto xb = static_cast<const __nv_bfloat16*>(x);
#define CASE(EP) router_dwg_partial_kernel<EP><<<grid, 256, 0, st>>>(dlogits, xb, T, d, E, tpc, scratch)
  if (E <= 8) CASE(8);
  else if (E <= 16) CASE(16);
  else if (E <= 32) CASE(32);
  else CASE(64);
#undef CASE
  if (int rc = check_launch("router_dwg_partial")) return rc;
  const int n = E * d;
  router_dwg_reduce_kernel<<<(n + 255) / 256, 256, 0, st>>>(scratch, chunks, n, dwg);
  return check_launch("router_dwg_reduce");
}

extern "C" int b200moe_router_bwd_dctx(const float* dlogits, int T, int E, int tokens_per_image, float* dctx_bias,
                                       void* stream) {
  B200MOE_REQUIRE(tokens_per_image > 0, B200MOE_EINVAL, "router_bwd_dctx: tokens_per_image <= 0");
  if (T == 0) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  const int B = (T + tokens_per_image - 1) / tokens_per_image;
  router_dctx_kernel<<<B, 256, 0, static_cast<cudaStream_t>(stream)>>>(dlogits, T, E, tokens_per_image, dctx_bias);
  return check_launch("router_dctx");
}
