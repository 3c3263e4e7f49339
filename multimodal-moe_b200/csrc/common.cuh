// Shared device helpers for libb200moe (sm_100a only): error plumbing, 128-bit global access,
// mbarrier / TMA / tcgen05 inline-PTX wrappers.  No torch types anywhere in csrc/.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "../../include/b200moe.h"

namespace b200moe {

// ---- host-side error plumbing (thread-local last-error string, see api.cu) ---------------------
void set_error(const char* fmt, ...);
int check_launch(const char* what);

#define B200MOE_REQUIRE(cond, code, ...)      \
  do {                                        \
    if (!(cond)) {                            \
      ::b200moe::set_error(__VA_ARGS__);      \
      return (code);                          \
    }                                         \
  } while (0)

// Programmatic dependent launch (PDL): inside the whole-layer calls every kernel is launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, so its blocks may be scheduled (and run their prologue:
// barrier init, TMEM allocation, descriptor prefetch) while the previous kernel of the stream drains.  Every kernel
// of the layer path executes pdl_sync() before its first global-memory access: griddepcontrol.wait returns only when
// the prerequisite grid has completed and flushed, so data dependencies are exactly those of plain stream order.
// Without the attribute both instructions are no-ops.
extern thread_local bool g_pdl;
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_sync() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = g_pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
// same, as clusters of two CTAs (grid.x even)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k_pair(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = g_pdl ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

// Multiprocessors of the current device (148 on a B200), read once per device (api.cu); 148 when no device is visible
// (the pure layout helpers of the ABI are callable on a CPU-only box).
int num_sms();
int current_device();          // cudaGetDevice; -1 on failure
// One-time, per-device kernel configuration (cudaFuncSetAttribute is a per-device attribute): `done` is false the first
// time a launcher runs on a device; two threads racing through it both set the same attribute, which is harmless.
struct DeviceOnce {
  std::atomic<unsigned long long> mask{0};
  bool done(int dev) const { return dev >= 0 && dev < 64 && ((mask.load(std::memory_order_acquire) >> dev) & 1ull); }
  void set(int dev) { if (dev >= 0 && dev < 64) mask.fetch_or(1ull << dev, std::memory_order_release); }
};
constexpr int kRowPad = 128;          // every expert segment of the permuted buffer starts at a multiple of this
constexpr int kRouterTokPerBlock = 128;  // max tokens per router/plan block (granularity of the priority scan)
// Tokens per router block: one 128-token tcgen05 M tile (router.cu).  Router, plan and the host wrappers all use this.
__host__ __device__ inline int router_block_tokens(int /*T*/) { return kRouterTokPerBlock; }

// ---- small device utilities -----------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ uint4 ldg128(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
// coherent at L2: for rows that other CTAs of the SAME kernel produced before a grid-wide rendezvous
__device__ __forceinline__ uint4 ldcg128(const void* p) {
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void stg128(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ float2 bf16x2_to_float2(uint32_t v) {
  // bf16 -> fp32 is a 16-bit shift
  return make_float2(__uint_as_float(v << 16), __uint_as_float(v & 0xffff0000u));
}
__device__ __forceinline__ uint32_t float2_to_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void unpack8(const uint4& r, float (&v)[8]) {
  float2 f;
  f = bf16x2_to_float2(r.x); v[0] = f.x; v[1] = f.y;
  f = bf16x2_to_float2(r.y); v[2] = f.x; v[3] = f.y;
  f = bf16x2_to_float2(r.z); v[4] = f.x; v[5] = f.y;
  f = bf16x2_to_float2(r.w); v[6] = f.x; v[7] = f.y;
}
__device__ __forceinline__ uint4 pack8(const float (&v)[8]) {
  uint4 r;
  r.x = float2_to_bf16x2(v[0], v[1]); r.y = float2_to_bf16x2(v[2], v[3]);
  r.z = float2_to_bf16x2(v[4], v[5]); r.w = float2_to_bf16x2(v[6], v[7]);
  return r;
}

// Context path of A.2 as the kernels see it (device pointers by value): ctx_bias[b,e] = sum_i emb_b[i] * wc[e, i] with
// emb_b the concatenation of the rows Embed_f[context[b,f]].  ctx_bias_dot is THE arithmetic of that dot product -- the
// stand-alone kernel (layer.cu) and the router's in-kernel evaluation both call it, so they produce the same bits.
struct CtxTables {
  const float* table[B200MOE_MAX_CTX_FIELDS];
  float* dtable[B200MOE_MAX_CTX_FIELDS];
  int vocab[B200MOE_MAX_CTX_FIELDS];
};
struct CtxArgs {
  const int64_t* context;   // [B,n] or null
  CtxTables tb;
  const float* wc;          // [E, n*cd]
  int n, cd;
};
#ifdef __CUDACC__
__device__ __forceinline__ float ctx_emb_value(const CtxArgs& c, int b, int i) {
  const int f = i / c.cd, j = i - f * c.cd;
  const long long id = c.context[static_cast<size_t>(b) * c.n + f];
  return (id >= 0 && id < c.tb.vocab[f]) ? c.tb.table[f][static_cast<size_t>(id) * c.cd + j] : 0.f;
}
// one warp: lanes stride the n*cd products in order, then the xor-shuffle tree (fixed order: deterministic)
template <typename EmbFn>
__device__ __forceinline__ float ctx_bias_dot(EmbFn emb, const float* __restrict__ wc_row, int ncd, int lane) {
  float s = 0.f;
  for (int i = lane; i < ncd; i += 32) s = fmaf(emb(i), wc_row[i], s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  return s;
}
#endif

// The A.7 combine y[t] = sum_s g_s * yo_s, fp32, slot order: product and sum are rounded separately (no FMA contraction)
// and flush subnormals, which is exactly what an fp32 reduction at L2 (red.global.add.f32) does with a pre-multiplied
// term.  A sum of k <= 2 such terms does not depend on their order, so the down-projection's fused epilogue (terms
// arrive in tile order) and the separate combine kernels (slot order) produce the same bits.
__device__ __forceinline__ float combine_term(float g, float v) {
  float r;
  asm("mul.rn.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(g), "f"(v));
  return r;
}
__device__ __forceinline__ float combine_add(float acc, float term) {
  float r;
  asm("add.rn.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(acc), "f"(term));
  return r;
}

// Rows that live in a peer GPU's memory (expert parallelism over NVLink, ep_p2p.cu): every rank maps the arenas of all
// ranks (symmetric allocation); a row is named by dest = rank * rows_cap + row inside that rank's buffer.
struct PeerRows {
  const unsigned char* const* peers;   // device array [W] of arena base addresses (as mapped on this GPU); NULL = local
  long long off_bytes;                 // offset of the buffer inside every arena
  const int32_t* dest_of_row;          // [send rows] -> rank * rows_cap + row
  int rows_cap;                        // rows of the buffer on every rank
};
__device__ __forceinline__ __nv_bfloat16* peer_row(const PeerRows& pr, int dest, int d) {
  const int w = dest / pr.rows_cap, row = dest - w * pr.rows_cap;
  return reinterpret_cast<__nv_bfloat16*>(const_cast<unsigned char*>(pr.peers[w]) + pr.off_bytes) +
         static_cast<size_t>(row) * d;
}

// erf-GELU (hybrid_encoder.py:157 with utils.py:160-161 -> nn.GELU()) for the GEMM epilogues.
// The epilogue has ~16 issue slots per output element before it, not the tensor pipe, bounds the up-projection
// (DESIGN.md), and erff() alone costs ~25.  Phi(x) = 0.5 (1 + erf(x / sqrt 2)) is evaluated as
// 0.5 (1 + tanh(u)), u = x (c0 + c1 x^2 + c2 x^4), a minimax fit of atanh(erf(x / sqrt 2)) (|dPhi| <= 1.9e-5 on
// [-8, 8]), with tanh on the MUFU pipe (tanh.approx.f32, rel 2^-11 => |dPhi| <= 2.5e-4):
// below the 2^-9 rounding of the bf16 outputs.  The derivative is the exact derivative of that same function, so
// forward and backward stay consistent:  gelu'(x) = Phi + x * 0.5 (1 - t^2) u'(x),  u' = c0 + 3 c1 x^2 + 5 c2 x^4.
constexpr float kGeluC0 = 0.7977178338960558f;
constexpr float kGeluC1 = 0.036798255994169915f;
constexpr float kGeluC2 = -0.0003158070465750329f;
__device__ __forceinline__ float tanh_mufu(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// |x| > 8 is handled by clamping x^2 only: the polynomial factor then stays at its value for |x| = 8 (1.86), u = 1.86 x
// has |u| > 14.9, tanh.approx returns exactly +-1, so Phi is exactly 0 or 1 and sech^2 exactly 0 (one FMNMX instead of two
// per element in kernels whose epilogue issue slots are the limit).
__device__ __forceinline__ float gelu_erf(float v) {
  const float x2 = fminf(v * v, 64.0f);
  const float t = tanh_mufu(v * fmaf(x2, fmaf(x2, kGeluC2, kGeluC1), kGeluC0));
  const float hv = 0.5f * v;
  return fmaf(hv, t, hv);
}
__device__ __forceinline__ float gelu_erf_grad(float v) {
  const float x2 = fminf(v * v, 64.0f);
  const float t = tanh_mufu(v * fmaf(x2, fmaf(x2, kGeluC2, kGeluC1), kGeluC0));
  const float up = fmaf(x2, fmaf(x2, 5.0f * kGeluC2, 3.0f * kGeluC1), kGeluC0);
  const float cdf = fmaf(0.5f, t, 0.5f);
  const float sech2 = fmaf(-t, t, 1.0f);
  return fmaf(0.5f * v * sech2, up, cdf);
}

// ---- mbarrier -----------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// ---- TMA ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 2-D tiled load, completion signalled as transaction bytes on `bar`.  c0 = innermost coordinate.
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               :: "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
               :: "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- tcgen05 / TMEM ------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot) {  // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
               :: "r"(smem_u32(smem_slot)), "r"(kCols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {  // same warp that allocated
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(taddr), "r"(kCols) : "memory");
}

// ---- CTA pairs (cta_group::2): two SMs of a cluster of 2 share one M=256 MMA tile ------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the mbarrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t rank) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}"
      :: "r"(smem_u32(bar)), "r"(rank) : "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc2(uint32_t* smem_slot) {  // one full warp, same warp id in both CTAs
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;"
               :: "r"(smem_u32(smem_slot)), "r"(kCols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" :: "r"(taddr), "r"(kCols) : "memory");
}
// 2-D tiled load issued by either CTA of a pair into its OWN shared memory; the transaction bytes are counted on the
// mbarrier of CTA 0 (peer bit of the shared::cluster address cleared), which is the one the MMA issuer waits on.
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
// Four rows of a 2-D tensor (map with a [1 x box] box) picked by index, laid down as four consecutive 128-byte-swizzled
// rows: the shared-memory image is that of rows r .. r+3 of a tiled box when dst = box + 128 r (the swizzle is a
// function of the address), negative or past-the-end rows arrive as zeros and still count towards complete_tx
// (tools/gather4_probe.cu, measured on sm_100a).
__device__ __forceinline__ void tma_gather4(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, const int4& rows) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, "
      "%7}], [%2];"
      :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(rows.x),
         "r"(rows.y), "r"(rows.z), "r"(rows.w)
      : "memory");
}
__device__ __forceinline__ void tma_gather4_pair(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0,
                                                 const int4& rows) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, "
      "{%3, %4, %5, %6, %7}], [%2];"
      :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0),
         "r"(rows.x), "r"(rows.y), "r"(rows.z), "r"(rows.w)
      : "memory");
}
// D[tmem of both CTAs] (+)= A[256 rows: 128 per CTA] * B[N: N/2 per CTA]; issued by ONE thread of CTA 0.
__device__ __forceinline__ void umma_bf16_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                  uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// Arrive on the mbarrier at this offset in BOTH CTAs once every tcgen05.mma issued so far has completed.
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               :: "r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3)) : "memory");
}

// D[tmem] (+)= A[smem] * B[smem]; bf16 inputs, fp32 accumulate; issued by ONE thread.
__device__ __forceinline__ void umma_bf16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// Arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               :: "r"(smem_u32(bar)) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread i of the warp receives lane (base_lane + i).
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor for tcgen05.mma (SWIZZLE_128B, sm_100 "version 1").
//   K-major  operand: rows of 64 bf16 (128 B); 8-row groups SBO apart; LBO unused.
//   MN-major operand: k-rows of 64 bf16 (128 B) along MN; 8-k-row groups SBO apart; 64-element MN chunks LBO apart.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= 1ull << 46;   // descriptor version (Blackwell)
  d |= 2ull << 61;   // SWIZZLE_128B
  return d;
}
// Instruction descriptor, kind::f16: bf16 x bf16 -> fp32.
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4)                       // D format: f32
         | (1u << 7)                     // A format: bf16
         | (1u << 10)                    // B format: bf16
         | ((a_mn_major ? 1u : 0u) << 15)
         | ((b_mn_major ? 1u : 0u) << 16)
         | (static_cast<uint32_t>(N >> 3) << 17)
         | (static_cast<uint32_t>(M >> 4) << 24);
}

}  // namespace b200moe
