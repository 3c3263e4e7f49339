// Router gradient of ONE token by ONE warp, shared by the stand-alone dlogits kernel
// (router_bwd.cu) and by the backward of the combine (movement.cu), whose last-arriving row of a token finishes the
// token's dlogits in the same launch -- both call this function, so they produce the same bits.
//   dL/dlogits[t,:] has three sources, all evaluated in fp32 from the saved logits:
//     gates : g_s = p_s / (S + eps), S = sum of the selected p  ->  dL/dp_{i_s} = (dg_s - sum_j dg_j g_j) / (S + eps)
//             (raw p when normalize_topk == 0); dropped slots have dg = 0 but stay in the normaliser
//     L_lb  : E * sum_e f_e P_e, f constant                      ->  dL/dp[t,e] += d_lb * E * f_e / T
//     L_z   : mean_t lse^2                                        ->  dL/dlogit[t,e] += d_z * 2 lse_t p[t,e] / T
//   followed by the softmax Jacobian  dlogit_e = p_e (dp_e - sum_j p_j dp_j).
// Lane l owns experts l and l + 32 (E <= 64); reductions are xor-shuffle trees (fixed order: deterministic).
#pragma once
#include <math.h>

#include "common.cuh"

namespace b200moe {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// dgate_coherent: dgate_row was written earlier in the SAME kernel by other warps (read at L2).
__device__ __forceinline__ void dlogits_token_warp(const float* __restrict__ logits, const int32_t* __restrict__ topk_idx,
                                                   const int32_t* __restrict__ row_of_assignment,
                                                   const float* dgate_row, bool dgate_coherent,
                                                   const int32_t* __restrict__ tokens_per_expert, float d_lb, float d_z,
                                                   int t, int T, int E, int k, int normalize, int lane,
                                                   float* __restrict__ dlogits) {
  const int e0 = lane, e1 = lane + 32;
  const float* lrow = logits + static_cast<size_t>(t) * E;
  const float l0 = e0 < E ? lrow[e0] : -INFINITY, l1 = e1 < E ? lrow[e1] : -INFINITY;
  const float mx = warp_max(fmaxf(l0, l1));
  float p0 = e0 < E ? expf(l0 - mx) : 0.f, p1 = e1 < E ? expf(l1 - mx) : 0.f;
  const float ssum = warp_sum(p0 + p1);
  const float inv = 1.0f / ssum;
  const float lse = mx + logf(ssum);
  p0 *= inv;
  p1 *= inv;
  const float invT = 1.0f / static_cast<float>(T);
  const float tk = static_cast<float>(static_cast<long long>(T) * k);
  float dp0 = e0 < E ? d_lb * static_cast<float>(E) * (static_cast<float>(tokens_per_expert[e0]) / tk) * invT : 0.f;
  float dp1 = e1 < E ? d_lb * static_cast<float>(E) * (static_cast<float>(tokens_per_expert[e1]) / tk) * invT : 0.f;
  if (dgate_row != nullptr) {           // gate path (every lane computes the same scalars)
    int idx[8];
    float dg[8], ps[8];
    float S = 0.f;
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      if (s < k) {
        idx[s] = topk_idx[static_cast<size_t>(t) * k + s];
        const int r = row_of_assignment[static_cast<size_t>(t) * k + s];
        dg[s] = r < 0 ? 0.f : (dgate_coherent ? __ldcg(dgate_row + r) : dgate_row[r]);
        const float a = __shfl_sync(0xffffffffu, p0, idx[s] & 31), b = __shfl_sync(0xffffffffu, p1, idx[s] & 31);
        ps[s] = idx[s] < 32 ? a : b;
        S += ps[s];
      }
    }
    const float den = S + 1e-9f;
    float corr = 0.f;
#pragma unroll
    for (int s = 0; s < 8; ++s)
      if (s < k && normalize) corr += dg[s] * (ps[s] / den);
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      if (s < k) {
        const float g = normalize ? (dg[s] - corr) / den : dg[s];
        dp0 += (idx[s] == e0) ? g : 0.f;
        dp1 += (idx[s] == e1) ? g : 0.f;
      }
    }
  }
  const float dot = warp_sum(fmaf(p0, dp0, p1 * dp1));
  const float zc = d_z * 2.0f * lse * invT;
  float* orow = dlogits + static_cast<size_t>(t) * E;
  if (e0 < E) orow[e0] = p0 * (dp0 - dot) + zc * p0;
  if (e1 < E) orow[e1] = p1 * (dp1 - dot) + zc * p1;
}

}  // namespace b200moe
