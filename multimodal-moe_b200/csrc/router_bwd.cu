// Router gradient.
//   dL/dlogits[t,:] has three sources, all evaluated in fp32 from the saved logits:
//     gates : g_s = p_s / (S + eps), S = sum of the selected p  ->  dL/dp_{i_s} = (dg_s - sum_j dg_j g_j) / (S + eps)
//             (raw p when normalize_topk == 0); dropped slots have dg = 0 but stay in the normaliser
//     L_lb  : E * sum_e f_e P_e, f constant                      ->  dL/dp[t,e] += d_lb * E * f_e / T
//     L_z   : mean_t lse^2                                        ->  dL/dlogit[t,e] += d_z * 2 lse_t p[t,e] / T
//   followed by the softmax Jacobian  dlogit_e = p_e (dp_e - sum_j p_j dp_j).
// Then dWg = dlogits^T x (two-stage fixed-order reduction), dctx_bias = per-image segment sums.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "router_grad.cuh"

namespace b200moe {

// one warp per token (router_grad.cuh).  128-thread blocks and no shared memory: at the small problem sizes this kernel
// runs on a side stream next to dgrad_down, whose persistent CTAs (576 threads x 96 registers, ~all of the shared memory)
// leave ~9 K registers per SM -- a 128-thread block fits beside them, a 256-thread block waited for an SM (round 2
// timeline at 3200 tokens: launched at 61 us, started at 71 us).
__global__ void __launch_bounds__(128)
router_dlogits_kernel(const float* __restrict__ logits, const int32_t* __restrict__ topk_idx,
                      const int32_t* __restrict__ row_of_assignment, const float* __restrict__ dgate_row,
                      const int32_t* __restrict__ tokens_per_expert, const float* __restrict__ d_aux, int T, int E,
                      int k, int normalize, float aux_w, float z_w, float* __restrict__ dlogits) {
  pdl_sync();
  const float da = d_aux ? *d_aux : 0.f;
  const int lane = threadIdx.x & 31;
  for (int t = blockIdx.x * 4 + (threadIdx.x >> 5); t < T; t += gridDim.x * 4)
    dlogits_token_warp(logits, topk_idx, row_of_assignment, dgate_row, false, tokens_per_expert, da * aux_w, da * z_w, t, T,
                       E, k, normalize, lane, dlogits);
}

// dWg = dlogits^T x in two fixed-order stages (deterministic, no atomics).
// stage 1: block = (token chunk, 256-column slab, group of 16 experts), 4 warps.  A lane owns two 4-column pieces of
//   the slab (c, c+128) and a [16 experts][8 columns] register tile; the warps take every 4th token of the chunk, the
//   chunk's dlogits pass through shared memory in batches of 64 tokens (broadcast reads), so one 16-byte read of x
//   feeds 128 FMAs.  The four warps' tiles are then summed in order through shared memory -> scratch[chunk][e][col].
// stage 2: sum over the chunks; the same launch also reduces dctx_bias[b,e] = sum of dlogits over the tokens of image b.
constexpr int kDwgEG = 16;
constexpr int kDwgBatch = 64;
constexpr int kDwgRedBytes = 4 * kDwgEG * 256 * 4;

__global__ void __launch_bounds__(128)
router_dwg_partial_kernel(const float* __restrict__ dlogits, const __nv_bfloat16* __restrict__ x, int T, int d, int E,
                          int tok_per_chunk, float* __restrict__ scratch) {
  extern __shared__ __align__(16) float red_s[];      // [4][16][256]
  __shared__ __align__(16) float dl_s[kDwgBatch][kDwgEG];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_sync();
  const int slab0 = blockIdx.y * 256, e0 = blockIdx.z * kDwgEG;
  const int c0 = slab0 + lane * 4, c1 = c0 + 128;
  const bool has0 = c0 < d, has1 = c1 < d;
  const int t_begin = blockIdx.x * tok_per_chunk;
  const int t_end = min(T, t_begin + tok_per_chunk);
  float acc[kDwgEG][8];
#pragma unroll
  for (int e = 0; e < kDwgEG; ++e)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[e][j] = 0.f;

  for (int tb = t_begin; tb < t_end; tb += kDwgBatch) {
    const int nt = min(kDwgBatch, t_end - tb);
    __syncthreads();
    for (int i = threadIdx.x; i < kDwgBatch * kDwgEG; i += 128) {
      const int tt = i / kDwgEG, e = e0 + (i % kDwgEG);
      dl_s[tt][i % kDwgEG] = (tt < nt && e < E) ? dlogits[static_cast<size_t>(tb + tt) * E + e] : 0.f;
    }
    __syncthreads();
#pragma unroll 4
    for (int i = warp; i < nt; i += 4) {
      const __nv_bfloat16* xr = x + static_cast<size_t>(tb + i) * d;
      const uint2 ra = has0 ? __ldg(reinterpret_cast<const uint2*>(xr + c0)) : make_uint2(0, 0);
      const uint2 rb = has1 ? __ldg(reinterpret_cast<const uint2*>(xr + c1)) : make_uint2(0, 0);
      float xv[8];
      float2 f;
      f = bf16x2_to_float2(ra.x); xv[0] = f.x; xv[1] = f.y;
      f = bf16x2_to_float2(ra.y); xv[2] = f.x; xv[3] = f.y;
      f = bf16x2_to_float2(rb.x); xv[4] = f.x; xv[5] = f.y;
      f = bf16x2_to_float2(rb.y); xv[6] = f.x; xv[7] = f.y;
#pragma unroll
      for (int e4 = 0; e4 < kDwgEG; e4 += 4) {
        const float4 g = *reinterpret_cast<const float4*>(&dl_s[i][e4]);
        const float gv[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[e4 + q][j] = fmaf(gv[q], xv[j], acc[e4 + q][j]);
      }
    }
  }
  // warps in fixed order
#pragma unroll
  for (int e = 0; e < kDwgEG; ++e) {
    float* r = red_s + (warp * kDwgEG + e) * 256;
    *reinterpret_cast<float4*>(r + lane * 4) = make_float4(acc[e][0], acc[e][1], acc[e][2], acc[e][3]);
    *reinterpret_cast<float4*>(r + 128 + lane * 4) = make_float4(acc[e][4], acc[e][5], acc[e][6], acc[e][7]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kDwgEG * 256; i += 128) {
    const int e = e0 + (i >> 8), col = slab0 + (i & 255);
    if (e < E && col < d) {
      const float v = ((red_s[i] + red_s[kDwgEG * 256 + i]) + red_s[2 * kDwgEG * 256 + i]) + red_s[3 * kDwgEG * 256 + i];
      scratch[(static_cast<size_t>(blockIdx.x) * E + e) * d + col] = v;
    }
  }
}

// blocks [0, nred): dwg[i] = sum over chunks of scratch[c][i];  blocks [nred, nred + B): dctx[b,:] for image b
// tensor-pipe partials (lo_off != 0): chunk c holds [128][d] = hi rows 0..63, lo rows 64..127 -> element i of dwg is
// scratch[c*cstride + i] + scratch[c*cstride + lo_off + i]
__global__ void __launch_bounds__(256)
router_dwg_reduce_kernel(const float* __restrict__ scratch, int nchunks, int n, float* __restrict__ dwg, int nred,
                         const float* __restrict__ dlogits, int T, int E, int tpi, float* __restrict__ dctx,
                         int64_t cstride, int lo_off) {
  pdl_sync();
  if (static_cast<int>(blockIdx.x) < nred) {
    // 32 outputs x 8 phases per block: phase p sums chunks p, p+8, ... (eight loads in flight), then the phases are
    // added in order -- a fixed summation order with 8x the memory parallelism of one thread per output
    __shared__ float ph_s[8][33];
    const int o = threadIdx.x & 31, p = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + o;
    float s = 0.f;
    if (i < n) {
      int c = p;
      for (; c + 56 < nchunks; c += 64) {
        float a[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const float* q = scratch + static_cast<size_t>(c + 8 * u) * cstride + i;
          a[u] = lo_off ? q[0] + q[lo_off] : q[0];
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) s += a[u];
      }
      for (; c < nchunks; c += 8) {
        const float* q = scratch + static_cast<size_t>(c) * cstride + i;
        s += lo_off ? q[0] + q[lo_off] : q[0];
      }
    }
    ph_s[p][o] = s;
    __syncthreads();
    if (p == 0 && i < n) {
      float r = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) r += ph_s[q][o];
      dwg[i] = r;
    }
    return;
  }
  // dctx: thread (phase, e) strides the image's tokens; consecutive threads read consecutive addresses
  __shared__ float part[256];
  const int b = blockIdx.x - nred;
  const int nphase = 256 / E;
  const int ph = threadIdx.x / E, e = threadIdx.x - ph * E;
  const int t0 = b * tpi, t1 = min(T, t0 + tpi);
  float s = 0.f;
  if (ph < nphase)
    for (int t = t0 + ph; t < t1; t += nphase) s += dlogits[static_cast<size_t>(t) * E + e];
  part[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x < E) {
    float r = 0.f;
    for (int q = 0; q < nphase; ++q) r += part[q * E + threadIdx.x];
    dctx[static_cast<size_t>(b) * E + threadIdx.x] = r;
  }
}

// ---- the same contraction on the tensor pipe (T >= kDwgTensorMinT) --------------------------------------------------
// dWg = dlogits^T x is a [E x T] x [T x d] product: 2 T E d FMAs, 84 us on CUDA cores at E = 64, d = 1024, T = 16384.  The
// weight-gradient mode of the grouped GEMM (gemm.cu) computes it with fake experts = chunks of rows: A = the bf16
// (hi | lo) split of dlogits ([T,128]: hi in columns 0..E-1, lo = bf16(v - hi) in columns 64..64+E-1 -- 16 mantissa bits,
// x is bf16 already), B = x, partial[c] [128,d] fp32, then the fixed-order reduction above adds chunks and hi + lo.
constexpr int kDwgTensorMinTDefault = 8192;
// B200MOE_DWG_TENSOR_MIN_T: measurement aid (a huge value keeps the CUDA-core path at every size)
static int dwg_tensor_min_t() {
  static const int v = [] { const char* e = getenv("B200MOE_DWG_TENSOR_MIN_T"); return e ? atoi(e) : kDwgTensorMinTDefault; }();
  return v;
}
#define kDwgTensorMinT dwg_tensor_min_t()
struct DwgPlan { int t64, chunk_rows, nchunks; };
static DwgPlan dwg_tensor_plan(int T, int d) {
  DwgPlan p;
  p.t64 = (T + 63) / 64 * 64;
  const int want = (2 * num_sms() + d / 128 - 1) / (d / 128);           // ~two tiles per SM
  p.chunk_rows = ((p.t64 + want - 1) / want + 63) / 64 * 64;
  p.nchunks = (p.t64 + p.chunk_rows - 1) / p.chunk_rows;
  return p;
}
// scratch layout (floats): partial [nchunks][128][d] | dl_split [t64][128] bf16 | chunk_offsets [nchunks + 1] int32
static int64_t dwg_tensor_floats(int T, int d) {
  const DwgPlan p = dwg_tensor_plan(T, d);
  return static_cast<int64_t>(p.nchunks) * 128 * d + static_cast<int64_t>(p.t64) * 64 + (p.nchunks + 1 + 63) / 64 * 64;
}

__global__ void __launch_bounds__(256)
router_dl_split_kernel(const float* __restrict__ dlogits, int T, int E, int t64, int chunk_rows, int nchunks,
                       __nv_bfloat16* __restrict__ split, int32_t* __restrict__ chunk_offsets) {
  pdl_sync();
  if (blockIdx.x == 0)
    for (int c = threadIdx.x; c <= nchunks; c += blockDim.x) chunk_offsets[c] = min(c * chunk_rows, t64);
  // a thread writes 8 of the 128 columns of a row (16 bytes)
  const int64_t n = static_cast<int64_t>(t64) * 16;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(i >> 4), c8 = static_cast<int>(i & 15) * 8;
    const bool lo = c8 >= 64;
    const int e0 = lo ? c8 - 64 : c8;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float x = (t < T && e0 + j < E) ? dlogits[static_cast<size_t>(t) * E + e0 + j] : 0.f;
      v[j] = lo ? x - __bfloat162float(__float2bfloat16_rn(x)) : x;
    }
    stg128(split + static_cast<size_t>(t) * 128 + c8, pack8(v));
  }
}

int launch_dwg_gemm(const void* dl_split, const void* x, int T, int N, const int32_t* chunk_offsets, int nchunks,
                    float* partial, cudaStream_t st);       // gemm_host.inc

static int dwg_chunks(int T, int d, int E) {
  const int slabs = (d + 255) / 256, groups = (E + kDwgEG - 1) / kDwgEG;
  int chunks = (4 * num_sms() + slabs * groups - 1) / (slabs * groups);
  const int max_chunks = (T + 15) / 16;
  if (chunks > max_chunks) chunks = max_chunks;
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
  int grid = (T + 3) / 4;
  if (grid > 128 * num_sms()) grid = 128 * num_sms();   // one token per warp up to 75 k tokens: every gather in flight at once
  launch_k(router_dlogits_kernel, grid, 128, 0, st, logits, topk_idx, row_of_assignment, dgate_row, tokens_per_expert,
           d_aux, T, E, k, normalize_topk, aux_loss_weight, z_loss_weight, dlogits);
  return check_launch("router_dlogits");
}

extern "C" int64_t b200moe_dwg_scratch_floats(int T, int d, int E) {
  const int64_t simt = static_cast<int64_t>(dwg_chunks(T, d, E)) * E * d;
  if (T < kDwgTensorMinT || d % 128 != 0) return simt;
  const int64_t tens = dwg_tensor_floats(T, d);
  return tens > simt ? tens : simt;
}

extern "C" int b200moe_router_bwd_params(const float* dlogits, const void* x, int T, int d, int E,
                                         int tokens_per_image, float* scratch, float* dwg, float* dctx_bias,
                                         void* stream) {
  B200MOE_REQUIRE(E >= 1 && E <= 64 && d > 0 && d % 128 == 0, B200MOE_EINVAL, "router_bwd_params: bad E=%d / d=%d", E, d);
  B200MOE_REQUIRE(dctx_bias == nullptr || tokens_per_image > 0, B200MOE_EINVAL, "router_bwd_params: tokens_per_image <= 0");
  B200MOE_REQUIRE(dwg == nullptr || (scratch != nullptr && x != nullptr), B200MOE_EINVAL, "router_bwd_params: NULL scratch / x");
  if (T == 0 || (dwg == nullptr && dctx_bias == nullptr)) return B200MOE_OK;
  if (int rc = b200moe_check_device()) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int chunks = 0;
  int64_t cstride = static_cast<int64_t>(E) * d;
  int lo_off = 0;
  if (dwg && T >= kDwgTensorMinT) {
    // tensor pipe: split -> grouped GEMM in weight-gradient mode over row chunks -> reduction (below)
    const DwgPlan p = dwg_tensor_plan(T, d);
    float* partial = scratch;
    __nv_bfloat16* split = reinterpret_cast<__nv_bfloat16*>(scratch + static_cast<int64_t>(p.nchunks) * 128 * d);
    int32_t* offs = reinterpret_cast<int32_t*>(scratch + static_cast<int64_t>(p.nchunks) * 128 * d +
                                               static_cast<int64_t>(p.t64) * 64);
    int blocks = static_cast<int>((static_cast<int64_t>(p.t64) * 16 + 255) / 256);
    if (blocks > 8 * num_sms()) blocks = 8 * num_sms();
    launch_k(router_dl_split_kernel, blocks, 256, 0, st, dlogits, T, E, p.t64, p.chunk_rows, p.nchunks, split, offs);
    if (int rc = check_launch("router_dl_split")) return rc;
    if (int rc = launch_dwg_gemm(split, x, T, d, offs, p.nchunks, partial, st)) return rc;
    chunks = p.nchunks;
    cstride = static_cast<int64_t>(128) * d;
    lo_off = 64 * d;
  } else if (dwg) {
    chunks = dwg_chunks(T, d, E);
    int tpc = (T + chunks - 1) / chunks;
    tpc = (tpc + 3) / 4 * 4;
    chunks = (T + tpc - 1) / tpc;
    static DeviceOnce once;
    const int dev = current_device();
    if (!once.done(dev)) {
      cudaError_t e = cudaFuncSetAttribute(router_dwg_partial_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDwgRedBytes);
      B200MOE_REQUIRE(e == cudaSuccess, B200MOE_ECUDA, "router_bwd_params: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      once.set(dev);
    }
    dim3 grid(chunks, (d + 255) / 256, (E + kDwgEG - 1) / kDwgEG);
    launch_k(router_dwg_partial_kernel, grid, 128, kDwgRedBytes, st, dlogits, static_cast<const __nv_bfloat16*>(x), T, d,
             E, tpc, scratch);
    if (int rc = check_launch("router_dwg_partial")) return rc;
  }
  const int n = E * d;
  const int nred = dwg ? (n + 31) / 32 : 0;
  const int B = dctx_bias ? (T + tokens_per_image - 1) / tokens_per_image : 0;
  launch_k(router_dwg_reduce_kernel, nred + B, 256, 0, st, scratch, chunks, n, dwg, nred, dlogits, T, E,
           tokens_per_image > 0 ? tokens_per_image : T, dctx_bias, cstride, lo_off);
  return check_launch("router_dwg_reduce");
}

extern "C" int b200moe_router_bwd_dwg(const float* dlogits, const void* x, int T, int d, int E, float* scratch,
                                      float* dwg, void* stream) {
  B200MOE_REQUIRE(dwg != nullptr, B200MOE_EINVAL, "router_bwd_dwg: dwg is NULL");
  return b200moe_router_bwd_params(dlogits, x, T, d, E, 0, scratch, dwg, nullptr, stream);
}

extern "C" int b200moe_router_bwd_dctx(const float* dlogits, int T, int E, int tokens_per_image, float* dctx_bias,
                                       void* stream) {
  B200MOE_REQUIRE(tokens_per_image > 0 && dctx_bias != nullptr, B200MOE_EINVAL, "router_bwd_dctx: bad arguments");
  return b200moe_router_bwd_params(dlogits, nullptr, T, 128, E, tokens_per_image, nullptr, nullptr, dctx_bias, stream);
}

// ---- b200moe_router_bwd_dwg without caller scratch -------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot allocate (code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The two-stage reduction's scratch
// lives in a file-scope device array of B200MOE_DWG_STATIC_MAX_FLOATS floats (16 MB; every GPU that loads the library
// holds it until exit).  In host code the array's NAME is a host-side stand-in, not a device address: the address is
// fetched once per device, with that device current, by cudaGetSymbolAddress and that pointer goes to the kernels.
// Every call on a device shares the array: calls that overlap on the GPU (different streams) race.  Nothing in it has
// to start at zero -- both paths (CUDA cores below 8192 tokens, tensor pipe from there on) write every element the
// reduction reads, the bf16 split and the chunk table included -- so a call is independent of the previous one's leftovers.
constexpr int64_t kDwgStaticFloats = B200MOE_DWG_STATIC_MAX_FLOATS;
static __device__ __align__(256) float g_dwg_static_scratch[kDwgStaticFloats];

extern "C" int b200moe_router_bwd_dwg_noalloc(const float* dlogits, const void* x, int T, int d, int E, float* dwg,
                                              void* stream) {
  B200MOE_REQUIRE(dlogits != nullptr && x != nullptr && dwg != nullptr, B200MOE_EINVAL, "router_bwd_dwg_noalloc: NULL argument");
  B200MOE_REQUIRE(T >= 0 && E >= 1 && E <= 64 && d > 0 && d % 128 == 0, B200MOE_EINVAL,
                  "router_bwd_dwg_noalloc: bad T=%d / E=%d / d=%d", T, E, d);
  const int64_t need = b200moe_dwg_scratch_floats(T, d, E);
  B200MOE_REQUIRE(need <= kDwgStaticFloats, B200MOE_EINVAL,
                  "router_bwd_dwg_noalloc: T=%d d=%d E=%d needs %lld scratch floats, the built-in array holds %lld "
                  "(B200MOE_DWG_STATIC_MAX_FLOATS); use b200moe_router_bwd_dwg with caller scratch",
                  T, d, E, static_cast<long long>(need), static_cast<long long>(kDwgStaticFloats));
  if (int rc = b200moe_check_device()) return rc;
  static DeviceOnce once;
  static float* addr_dev[64];
  const int dev = current_device();
  B200MOE_REQUIRE(dev >= 0 && dev < 64, B200MOE_ECUDA, "router_bwd_dwg_noalloc: no current CUDA device");
  if (!once.done(dev)) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_dwg_static_scratch);
    B200MOE_REQUIRE(e == cudaSuccess && p != nullptr, B200MOE_ECUDA, "router_bwd_dwg_noalloc: cudaGetSymbolAddress: %s",
                    cudaGetErrorString(e));
    addr_dev[dev] = p;
    once.set(dev);
  }
  return b200moe_router_bwd_params(dlogits, x, T, d, E, 0, addr_dev[dev], dwg, nullptr, stream);
}
