// Fused router, forward (SURVEY.md Appendix A.2-A.4 and the statistics of A.8/A.9).
//
// One warp per token: the lanes split the d-long dot products with 128-bit loads of the bf16 token row, the gate
// matrix sits in shared memory in fp32, partial sums are all-reduced with xor shuffles so that every lane holds
// all E logits, then softmax / top-k / gate renormalisation run in registers in fp32.  Ranking is done on
// (logit desc, expert index asc): softmax is monotone, so this is the order A.4 asks for, without the ties that
// fp32 rounding of p could fabricate.  A block owns B200MOE_ROUTER_BLOCK consecutive tokens and emits the
// per-(slot, expert) assignment counts and the sum_t p[t,e] / sum_t lse^2 partials that the plan kernel scans
// in a fixed order (deterministic, no float atomics).
#include <math.h>

#include "common.cuh"

namespace b200moe {

constexpr int kRouterWarps = 8;
constexpr int kRouterThreads = kRouterWarps * 32;

template <int EP, bool W_IN_SMEM>
__global__ void __launch_bounds__(kRouterThreads)
router_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ ctx_bias, int tokens_per_image,
                  const float* __restrict__ wg, int T, int d, int E, int k, int normalize, int tb,
                  float* __restrict__ logits_out, int32_t* __restrict__ topk_idx, float* __restrict__ topk_gate,
                  int32_t* __restrict__ block_counts, float* __restrict__ block_stats) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* w_s = reinterpret_cast<float*>(smem_raw);                       // [E][d] when W_IN_SMEM
  float* stat_s = w_s + (W_IN_SMEM ? static_cast<size_t>(E) * d : 0);    // [warps][EP+1]
  int* cnt_s = reinterpret_cast<int*>(stat_s + kRouterWarps * (EP + 1));  // [k][E]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (W_IN_SMEM) {
    const int n4 = (E * d) >> 2;
    for (int i = threadIdx.x; i < n4; i += blockDim.x)
      reinterpret_cast<float4*>(w_s)[i] = reinterpret_cast<const float4*>(wg)[i];
  }
  for (int i = threadIdx.x; i < k * E; i += blockDim.x) cnt_s[i] = 0;
  __syncthreads();
  const float* wsrc = W_IN_SMEM ? w_s : wg;

  const int t0 = blockIdx.x * tb;
  const int nchunk = d >> 3;  // 16-byte chunks per row
  float psum0 = 0.f, psum1 = 0.f, zsum = 0.f;  // lane e owns sum_t p[t,e] (and e+32)

  for (int i = warp; i < tb; i += kRouterWarps) {
    const int t = t0 + i;
    if (t >= T) break;
    float acc[EP];
#pragma unroll
    for (int e = 0; e < EP; ++e) acc[e] = 0.f;
    const __nv_bfloat16* xr = x + static_cast<size_t>(t) * d;
    for (int c = lane; c < nchunk; c += 32) {
      const uint4 raw = ldg128(xr + c * 8);
      float xv[8];
      float2 f;
      f = bf16x2_to_float2(raw.x); xv[0] = f.x; xv[1] = f.y;
      f = bf16x2_to_float2(raw.y); xv[2] = f.x; xv[3] = f.y;
      f = bf16x2_to_float2(raw.z); xv[4] = f.x; xv[5] = f.y;
      f = bf16x2_to_float2(raw.w); xv[6] = f.x; xv[7] = f.y;
#pragma unroll
      for (int e = 0; e < EP; ++e) {
        if (e < E) {
          const float4 wa = *reinterpret_cast<const float4*>(wsrc + static_cast<size_t>(e) * d + c * 8);
          const float4 wb = *reinterpret_cast<const float4*>(wsrc + static_cast<size_t>(e) * d + c * 8 + 4);
          float a = acc[e];
          a = fmaf(xv[0], wa.x, a); a = fmaf(xv[1], wa.y, a); a = fmaf(xv[2], wa.z, a); a = fmaf(xv[3], wa.w, a);
          a = fmaf(xv[4], wb.x, a); a = fmaf(xv[5], wb.y, a); a = fmaf(xv[6], wb.z, a); a = fmaf(xv[7], wb.w, a);
          acc[e] = a;
        }
      }
    }
    // all-reduce over the lanes, add the per-image context bias (A.2: additive on the logits)
    const float* cb = ctx_bias ? ctx_bias + static_cast<size_t>(t / tokens_per_image) * E : nullptr;
    float mx = -INFINITY;
#pragma unroll
    for (int e = 0; e < EP; ++e) {
      float v = warp_sum(acc[e]);
      if (e < E) {
        if (cb) v += __ldg(cb + e);
        mx = fmaxf(mx, v);
      } else {
        v = -INFINITY;
      }
      acc[e] = v;
    }
    float ssum = 0.f;
    float p[EP];
#pragma unroll
    for (int e = 0; e < EP; ++e) {
      p[e] = (e < E) ? expf(acc[e] - mx) : 0.f;
      ssum += p[e];
    }
    const float inv = 1.0f / ssum;
    const float lse = mx + logf(ssum);
    // top-k by (logit desc, index asc); every lane computes the same answer
    unsigned long long taken = 0ull;
    int my_idx = 0;       // lane s keeps slot s
    float my_gate = 0.f;
    float gsum = 0.f;
    for (int s = 0; s < k; ++s) {
      float best = -INFINITY;
      int bi = 0;
      bool found = false;
#pragma unroll
      for (int e = 0; e < EP; ++e) {
        if (e < E && !((taken >> e) & 1ull)) {
          if (!found || acc[e] > best) { best = acc[e]; bi = e; found = true; }
        }
      }
      taken |= 1ull << bi;
      float pb = 0.f;
#pragma unroll
      for (int e = 0; e < EP; ++e) pb = (e == bi) ? p[e] * inv : pb;
      gsum += pb;
      if (lane == s) { my_idx = bi; my_gate = pb; }
    }
    if (normalize) my_gate = my_gate / (gsum + 1e-9f);
    if (lane < k) {
      topk_idx[static_cast<size_t>(t) * k + lane] = my_idx;
      topk_gate[static_cast<size_t>(t) * k + lane] = my_gate;
      atomicAdd(&cnt_s[lane * E + my_idx], 1);
    }
    // logits out + per-lane statistics
    float lo = 0.f, hi = 0.f, plo = 0.f, phi = 0.f;
#pragma unroll
    for (int e = 0; e < EP; ++e) {
      if (e < 32) { if (lane == e) { lo = acc[e]; plo = p[e] * inv; } }
      else        { if (lane == e - 32) { hi = acc[e]; phi = p[e] * inv; } }
    }
    if (lane < E) logits_out[static_cast<size_t>(t) * E + lane] = lo;
    if (EP > 32 && lane + 32 < E) logits_out[static_cast<size_t>(t) * E + lane + 32] = hi;
    psum0 += plo;
    psum1 += phi;
    zsum += lse * lse;
  }
  // block partials in a fixed warp order
  if (lane < EP) stat_s[warp * (EP + 1) + lane] = psum0;
  if (EP > 32 && lane + 32 < EP) stat_s[warp * (EP + 1) + lane + 32] = psum1;
  if (lane == 0) stat_s[warp * (EP + 1) + EP] = zsum;
  __syncthreads();
  for (int e = threadIdx.x; e <= E; e += blockDim.x) {
    const int col = (e == E) ? EP : e;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kRouterWarps; ++w) s += stat_s[w * (EP + 1) + col];
    block_stats[static_cast<size_t>(blockIdx.x) * (E + 1) + e] = s;
  }
  for (int i = threadIdx.x; i < k * E; i += blockDim.x)
    block_counts[static_cast<size_t>(blockIdx.x) * k * E + i] = cnt_s[i];
}

template <int EP>
static int launch_router(const void* x, const float* ctx_bias, int tpi, const float* wg, int T, int d, int E, int k,
                         int normalize, float* logits, int32_t* idx, float* gate, int32_t* bc, float* bs,
                         cudaStream_t st) {
  const int tb = router_block_tokens(T);
  const int nblk = (T + tb - 1) / tb;
  const size_t tail = sizeof(float) * kRouterWarps * (EP + 1) + sizeof(int) * k * E;
  const size_t wbytes = sizeof(float) * static_cast<size_t>(E) * d;
  const bool in_smem = wbytes + tail <= 160 * 1024;
  const size_t smem = (in_smem ? wbytes : 0) + tail;
  auto xb = static_cast<const __nv_bfloat16*>(x);
  if (in_smem) {
    auto kern = router_fwd_kernel<EP, true>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    kern<<<nblk, kRouterThreads, smem, st>>>(xb, ctx_bias, tpi, wg, T, d, E, k, normalize, tb, logits, idx, gate, bc, bs);
  } else {
    router_fwd_kernel<EP, false><<<nblk, kRouterThreads, smem, st>>>(xb, ctx_bias, tpi, wg, T, d, E, k, normalize, tb,
                                                                     logits, idx, gate, bc, bs);
  }
  return check_launch("router_fwd");
}

}  // namespace b200moe

extern "C" int b200moe_router_fwd(const void* x, const float* ctx_bias, int tokens_per_image, const float* wg,
                                  int T, int d, int E, int k, int normalize_topk, float* logits, int32_t* topk_idx,
                                  float* topk_gate, int32_t* block_counts, float* block_stats, void* stream) {
  using namespace b200moe;
  B200MOE_REQUIRE(T >= 0 && d > 0 && d % 8 == 0, B200MOE_EINVAL, "router_fwd: d=%d must be a positive multiple of 8", d);
  B200MOE_REQUIRE(E >= 1 && E <= 64, B200MOE_EINVAL, "router_fwd: num_experts=%d outside [1,64]", E);
  B200MOE_REQUIRE(k >= 1 && k <= E && k <= 8, B200MOE_EINVAL, "router_fwd: top_k=%d must be in [1,min(E,8)]", k);
  B200MOE_REQUIRE(ctx_bias == nullptr || tokens_per_image > 0, B200MOE_EINVAL, "router_fwd: tokens_per_image <= 0");
  if (T == 0) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (tokens_per_image <= 0) tokens_per_image = T;
#define B200MOE_ROUTER_CASE(EP)                                                                              \
  return launch_router<EP>(x, ctx_bias, tokens_per_image, wg, T, d, E, k, normalize_topk, logits, topk_idx, \
                           topk_gate, block_counts, block_stats, st)
  if (E <= 8) B200MOE_ROUTER_CASE(8);
  if (E <= 16) B200MOE_ROUTER_CASE(16);
  if (E <= 32) B200MOE_ROUTER_CASE(32);
  B200MOE_ROUTER_CASE(64);
#undef B200MOE_ROUTER_CASE
}
