// Fused post-norm of the encoder layer (hybrid_encoder.py:158-160): out = LayerNorm(residual + combine), one warp per
// token.  Shared by the stand-alone kernels of movement.cu and by the tail of the down-projection (gemm.cu), which
// normalises straight out of the fp32 combine accumulator once every CTA has delivered its reductions.
#pragma once
#include "common.cuh"

namespace b200moe {

constexpr int kLnMaxI = 4;
// 8 consecutive elements of a [T,d] tensor that is bf16 or fp32 (the encoder's residual stream is fp32 under the
// reference's autocast: LayerNorm outputs fp32, hybrid_encoder.py:150-160 with det_engine.py:45)
__device__ __forceinline__ void load8_any(const void* base, bool f32, size_t elem, float (&a)[8]) {
  if (f32) {
    const float4 p = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(base) + elem));
    const float4 q = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(base) + elem) + 1);
    a[0] = p.x; a[1] = p.y; a[2] = p.z; a[3] = p.w; a[4] = q.x; a[5] = q.y; a[6] = q.z; a[7] = q.w;
  } else {
    unpack8(ldg128(static_cast<const __nv_bfloat16*>(base) + elem), a);
  }
}
__device__ __forceinline__ void store8_any(void* base, bool f32, size_t elem, const float (&o)[8]) {
  if (f32) {
    float* p = static_cast<float*>(base) + elem;
    *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(o[4], o[5], o[6], o[7]);
  } else {
    stg128(static_cast<__nv_bfloat16*>(base) + elem, pack8(o));
  }
}
template <int NI>
__device__ __forceinline__ void ln_gather_sum(const __nv_bfloat16* __restrict__ yo, const int32_t* __restrict__ roa,
                                              const float* __restrict__ gate, const float* __restrict__ y_acc,
                                              bool acc_cg, const void* __restrict__ res, bool res_f32,
                                              int t, int k, int d, int lane, float (&v)[NI][8]) {
  const int nchunk = d >> 3;
#pragma unroll
  for (int i = 0; i < NI; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) v[i][j] = 0.f;
  if (y_acc) {            // the combine already happened in the down-projection's epilogue (fp32, same arithmetic)
#pragma unroll
    for (int i = 0; i < NI; ++i) {
      const int c = lane + 32 * i;
      if (c < nchunk) {
        const float4* ap = reinterpret_cast<const float4*>(y_acc + static_cast<size_t>(t) * d + c * 8);
        const float4 a = acc_cg ? __ldcg(ap) : __ldg(ap);
        const float4 b = acc_cg ? __ldcg(ap + 1) : __ldg(ap + 1);
        v[i][0] = a.x; v[i][1] = a.y; v[i][2] = a.z; v[i][3] = a.w;
        v[i][4] = b.x; v[i][5] = b.y; v[i][6] = b.z; v[i][7] = b.w;
      }
    }
    k = 0;
  }
  for (int s = 0; s < k; ++s) {
    const int r = roa[static_cast<size_t>(t) * k + s];
    if (r < 0) continue;
    const float g = gate[static_cast<size_t>(t) * k + s];
#pragma unroll
    for (int i = 0; i < NI; ++i) {
      const int c = lane + 32 * i;
      if (c < nchunk) {
        float a[8];
        const __nv_bfloat16* src = yo + static_cast<size_t>(r) * d + c * 8;
        unpack8(acc_cg ? ldcg128(src) : ldg128(src), a);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] = combine_add(v[i][j], combine_term(g, a[j]));
      }
    }
  }
#pragma unroll
  for (int i = 0; i < NI; ++i) {
    const int c = lane + 32 * i;
    if (c < nchunk) {
      float a[8];
      load8_any(res, res_f32, static_cast<size_t>(t) * d + c * 8, a);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[i][j] += a[j];
    }
  }
}

// A.7 for one token, one warp: y[t] = sum_s gate[t,s] * yo[row[t,s]] in slot order, fp32 (combine_term / combine_add).
// `cg`: yo was written earlier in the SAME kernel by other CTAs (read it coherently at L2).
__device__ __forceinline__ void combine_token(const __nv_bfloat16* __restrict__ yo, const int32_t* __restrict__ roa,
                                              const float* __restrict__ gate, bool cg, int t, int k, int d, int lane,
                                              __nv_bfloat16* __restrict__ y) {
  const int nchunk = d >> 3;
  for (int c = lane; c < nchunk; c += 32) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int s = 0; s < k; ++s) {
      const int r = roa[static_cast<size_t>(t) * k + s];
      if (r < 0) continue;
      const float g = gate[static_cast<size_t>(t) * k + s];
      const __nv_bfloat16* src = yo + static_cast<size_t>(r) * d + c * 8;
      float v[8];
      unpack8(cg ? ldcg128(src) : ldg128(src), v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = combine_add(acc[j], combine_term(g, v[j]));
    }
    stg128(y + static_cast<size_t>(t) * d + c * 8, pack8(acc));
  }
}

// One token: gather / read the combine, add the residual, normalise, write out and (mean, rstd).  `acc_cg`: read the
// accumulator with ld.global.cg (it was produced by L2 reductions earlier in the SAME kernel).
template <int NI>
__device__ __forceinline__ void ln_forward_token(const __nv_bfloat16* __restrict__ yo, const int32_t* __restrict__ roa,
                                                 const float* __restrict__ gate, const float* __restrict__ y_acc,
                                                 bool acc_cg, const void* __restrict__ res, bool res_f32,
                                                 const float* __restrict__ gamma, const float* __restrict__ beta,
                                                 float eps, int t, int k, int d, int lane, void* __restrict__ out,
                                                 bool out_f32, float* __restrict__ mean_rstd) {
  const int nchunk = d >> 3;
  const float invd = 1.0f / static_cast<float>(d);
  float v[NI][8];
  ln_gather_sum<NI>(yo, roa, gate, y_acc, acc_cg, res, res_f32, t, k, d, lane, v);
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NI; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) s += v[i][j];          // missing chunks are zero
  const float mean = warp_sum(s) * invd;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NI; ++i)
    if (lane + 32 * i < nchunk)
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float c_ = v[i][j] - mean; q = fmaf(c_, c_, q); }
  const float rstd = rsqrtf(warp_sum(q) * invd + eps);
#pragma unroll
  for (int i = 0; i < NI; ++i) {
    const int c = lane + 32 * i;
    if (c < nchunk) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c * 8)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c * 8) + 1);
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + c * 8)), b1 = __ldg(reinterpret_cast<const float4*>(beta + c * 8) + 1);
      const float gm[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      const float bt[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = fmaf((v[i][j] - mean) * rstd, gm[j], bt[j]);
      store8_any(out, out_f32, static_cast<size_t>(t) * d + c * 8, o);
    }
  }
  if (lane == 0) { mean_rstd[2 * static_cast<size_t>(t)] = mean; mean_rstd[2 * static_cast<size_t>(t) + 1] = rstd; }
}

}  // namespace b200moe
