// Whole-layer entry points (include/b200moe.h): one C call runs every kernel of the forward (or backward) pass of
// the context-gated MoE layer on the caller's stream, out of two caller-owned workspaces, so the host language pays
// one FFI crossing per pass instead of one per kernel.  Also the context-bias kernels.
#include "common.cuh"

namespace b200moe {

static inline int64_t align256(int64_t v) { return (v + 255) & ~static_cast<int64_t>(255); }

// Optional per-kernel CUDA-event timing of the whole-layer calls (bench.py's roofline pass): events are recorded on
// the launching stream around every step and read back with b200moe_layer_profile_read after a synchronize.
constexpr int kMaxProf = 4096;
struct ProfEntry { const char* name; cudaEvent_t a, b; };
static bool g_prof_on = false;
static ProfEntry g_prof[kMaxProf];
static int g_prof_n = 0;
static bool g_pdl_enabled = true;

// The host needs a few microseconds between recording an event and getting the next kernel into the queue (tensor-map
// encoding, the launch itself); on an idle GPU that gap would be charged to the kernel.  A short spin kernel ahead of
// the first event keeps the GPU busy until the launch is queued, so the two events bracket device time only.
__global__ void prof_delay_kernel(long long ns) {
  long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do {
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  } while (t1 - t0 < ns);
}

struct ProfScope {
  cudaStream_t st;
  bool live;
  ProfScope(const char* name, void* stream) : st(static_cast<cudaStream_t>(stream)), live(g_prof_on && g_prof_n < kMaxProf) {
    if (!live) return;
    ProfEntry& e = g_prof[g_prof_n];
    e.name = name;
    cudaEventCreate(&e.a);
    cudaEventCreate(&e.b);
    prof_delay_kernel<<<1, 1, 0, st>>>(20000);
    cudaEventRecord(e.a, st);
  }
  ~ProfScope() {
    if (!live) return;
    cudaEventRecord(g_prof[g_prof_n].b, st);
    ++g_prof_n;
  }
};
// PDL on for the kernels launched inside one whole-layer call (off while the per-step profile is recording: the
// events between the kernels serialise them anyway and the timings should be those of isolated kernels)
struct PdlScope {
  bool prev;
  PdlScope() : prev(g_pdl) { g_pdl = !g_prof_on && g_pdl_enabled; }
  ~PdlScope() { g_pdl = prev; }
};
#define STEP(name, call)            \
  do {                              \
    ProfScope ps_(name, stream);    \
    rc = (call);                    \
  } while (0);                      \
  if (rc) return rc

// Backward-pass fork: the weight gradients / bias sums and the router's parameter gradients do not feed dx, so they
// run on two library-owned side streams, joined to the caller's stream before the call returns
// (plain event fork/join, so the pattern is captured into a CUDA graph as parallel branches).
struct ForkState {
  bool ok = false;
  cudaStream_t s0 = nullptr, s1 = nullptr, s2 = nullptr, s3 = nullptr;
  cudaEvent_t ev[6] = {};
};
static ForkState g_fork[16];
static int g_fork_mode = 1;   // 0 never, 1 always (measured faster at every size of the sweep: 3200 .. 65600 tokens)
static ForkState* fork_state() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
  ForkState& f = g_fork[dev];
  if (!f.ok) {
    // s0 carries ONLY the dlogits kernel, at the highest priority.  The hardware launches the CTAs of queued grids in order
    // within a priority level: dgrad_up, launched programmatically behind dgrad_down, sits in that queue with CTAs that
    // cannot be placed until dgrad_down's retire, and every grid behind it waits too (round-2 timelines: dlogits ready at
    // 58 us, started at 70 us at 3200 tokens; at C/65533 it started when dgrad_down ended and then ran 219 us beside
    // dgrad_up).  dlogits blocks are 128 threads without shared memory and fit beside a GEMM CTA; the rest of the router
    // branch uses shared memory and stays at normal priority (at the highest priority its blocks took the SMs the next
    // GEMM of the dx chain was waiting for: dgrad_up started 83 us late at C/65533).
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (cudaStreamCreateWithPriority(&f.s0, cudaStreamNonBlocking, prio_hi) != cudaSuccess) return nullptr;
    if (cudaStreamCreateWithFlags(&f.s1, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (cudaStreamCreateWithFlags(&f.s2, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (cudaStreamCreateWithFlags(&f.s3, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    for (auto& e : f.ev)
      if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    f.ok = true;
  }
  return &f;
}
// a kernel that directly follows a cross-stream wait is launched without the programmatic-serialization attribute
struct PdlOffIf {
  bool prev;
  explicit PdlOffIf(bool off) : prev(g_pdl) { if (off) g_pdl = false; }
  ~PdlOffIf() { g_pdl = prev; }
};

// combine inside the down-projection kernel (its tail, or -- reduction mode -- its epilogue: order-free fp32 accumulation
// is exact for at most two terms)
static inline bool fused_combine(const b200moe_layer_desc* q) {
  if (q->flags & B200MOE_LAYER_SEPARATE_COMBINE) return false;
  if (q->flags & (B200MOE_LAYER_COMBINE_TAIL | B200MOE_LAYER_COMBINE_REDUCE)) return true;
  return static_cast<int64_t>(q->rows_cap) * q->d * 2 <= (48ll << 20);     // yo still L2-resident when the tail reads it
}
static inline bool reduce_combine(const b200moe_layer_desc* q) {
  return fused_combine(q) && (q->flags & B200MOE_LAYER_COMBINE_REDUCE) && q->k <= 2;
}
// dispatch gather inside the GEMM operand loads (no permuted copy of x): opt-in, measured slower (include/b200moe.h)
static inline bool gather_dispatch(const b200moe_layer_desc* q) {
  return (q->flags & B200MOE_LAYER_GATHER_DISPATCH) && !reduce_combine(q);
}

static int check_desc(const b200moe_layer_desc* q, const char* what) {
  B200MOE_REQUIRE(q != nullptr, B200MOE_EINVAL, "%s: NULL descriptor", what);
  B200MOE_REQUIRE(q->T > 0, B200MOE_EINVAL, "%s: T=%d must be positive", what, q->T);
  B200MOE_REQUIRE(q->d >= 128 && q->d % 128 == 0 && q->F >= 128 && q->F % 128 == 0, B200MOE_EINVAL,
                  "%s: d_model=%d and dim_feedforward=%d must be multiples of 128", what, q->d, q->F);
  B200MOE_REQUIRE(q->E >= 1 && q->E <= 64 && q->k >= 1 && q->k <= 8 && q->k <= q->E, B200MOE_EINVAL,
                  "%s: bad num_experts=%d / top_k=%d", what, q->E, q->k);
  B200MOE_REQUIRE(q->capacity >= 0, B200MOE_EINVAL, "%s: negative capacity", what);
  B200MOE_REQUIRE(q->activation == B200MOE_ACT_GELU || q->activation == B200MOE_ACT_RELU, B200MOE_EINVAL,
                  "%s: activation=%d must be B200MOE_ACT_GELU or B200MOE_ACT_RELU", what, q->activation);
  B200MOE_REQUIRE((q->flags & ~(B200MOE_LAYER_SEPARATE_COMBINE | B200MOE_LAYER_COMBINE_REDUCE | B200MOE_LAYER_COMBINE_TAIL |
                                 B200MOE_LAYER_GATHER_DISPATCH)) == 0,
                  B200MOE_EINVAL,
                  "%s: unknown flags 0x%x", what, q->flags);
  B200MOE_REQUIRE(q->rows_cap > 0 && q->rows_cap % kRowPad == 0 &&
                      q->rows_cap >= b200moe_layer_rows_cap(q->T, q->k, q->E, q->capacity, q->d, q->F),
                  B200MOE_EINVAL, "%s: rows_cap=%d is below b200moe_layer_rows_cap(T,k,E,capacity,d,F)", what, q->rows_cap);
  return B200MOE_OK;
}

// ---- context bias ---------------------------------------------------------------------------------------------
// one block per image, thread per expert-strided output; the n*cd-long embedding row sits in shared memory
__global__ void __launch_bounds__(512)
ctx_bias_fwd_kernel(const int64_t* __restrict__ context, const CtxTables tb, int n, int cd,
                    const float* __restrict__ wc, int E, float* __restrict__ out) {
  extern __shared__ float emb_s[];   // [n*cd]
  const int b = blockIdx.x;
  for (int i = threadIdx.x; i < n * cd; i += blockDim.x) {
    const int f = i / cd, j = i - f * cd;
    const long long id = context[static_cast<size_t>(b) * n + f];
    emb_s[i] = (id >= 0 && id < tb.vocab[f]) ? tb.table[f][static_cast<size_t>(id) * cd + j] : 0.f;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int e = warp; e < E; e += (blockDim.x >> 5)) {
    const float s = ctx_bias_dot([&](int i) { return emb_s[i]; }, wc + static_cast<size_t>(e) * n * cd, n * cd, lane);
    if (lane == 0) out[static_cast<size_t>(b) * E + e] = s;
  }
}

// dwc[e, i] = sum_b dctx[b,e] * emb_b[i];  dtable_f[v, j] = sum_{b: ctx[b,f]=v} sum_e dctx[b,e] * wc[e, f*cd+j]
// One warp per output element: the lanes stride the images, then a fixed shuffle tree (deterministic).
__global__ void __launch_bounds__(256)
ctx_bias_bwd_kernel(const int64_t* __restrict__ context, const float* __restrict__ dctx, const CtxTables tb, int n,
                    int cd, const float* __restrict__ wc, int B, int E, float* __restrict__ dwc, int n_dwc,
                    int n_tab) {
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i < n_dwc) {
    const int e = i / (n * cd), c = i - e * (n * cd);
    const int f = c / cd, j = c - f * cd;
    float s = 0.f;
    for (int b = lane; b < B; b += 32) {
      const long long id = context[static_cast<size_t>(b) * n + f];
      if (id >= 0 && id < tb.vocab[f]) s = fmaf(dctx[static_cast<size_t>(b) * E + e], tb.table[f][static_cast<size_t>(id) * cd + j], s);
    }
    s = warp_sum(s);
    if (lane == 0) dwc[i] = s;
  } else if (i < n_dwc + n_tab) {
    int r = i - n_dwc;
    int f = 0;
    while (r >= tb.vocab[f] * cd) { r -= tb.vocab[f] * cd; ++f; }
    const int v = r / cd, j = r - v * cd;
    float s = 0.f;
    for (int b = lane; b < B; b += 32) {
      if (context[static_cast<size_t>(b) * n + f] != v) continue;
      float g = 0.f;
      for (int e = 0; e < E; ++e) g = fmaf(dctx[static_cast<size_t>(b) * E + e], wc[static_cast<size_t>(e) * n * cd + f * cd + j], g);
      s += g;
    }
    s = warp_sum(s);
    if (lane == 0) tb.dtable[f][r] = s;
  }
}

static int fill_tables(CtxTables* tb, const float* const* tables, float* const* dtables, const int32_t* vocab, int n,
                       const char* what) {
  B200MOE_REQUIRE(n >= 1 && n <= B200MOE_MAX_CTX_FIELDS, B200MOE_EINVAL, "%s: n_fields=%d outside [1,%d]", what, n,
                  B200MOE_MAX_CTX_FIELDS);
  for (int i = 0; i < B200MOE_MAX_CTX_FIELDS; ++i) {
    tb->table[i] = i < n ? tables[i] : nullptr;
    tb->dtable[i] = (i < n && dtables) ? dtables[i] : nullptr;
    tb->vocab[i] = i < n ? vocab[i] : 0;
    if (i < n) B200MOE_REQUIRE(tables[i] != nullptr && vocab[i] > 0, B200MOE_EINVAL, "%s: bad table %d", what, i);
  }
  return B200MOE_OK;
}

}  // namespace b200moe

using namespace b200moe;

extern "C" int b200moe_ctx_bias_fwd(const int64_t* context, const float* const* tables, const int32_t* vocab,
                                    int n_fields, int context_dim, const float* wc, int B, int E, float* ctx_bias,
                                    void* stream) {
  CtxTables tb;
  if (int rc = fill_tables(&tb, tables, nullptr, vocab, n_fields, "ctx_bias_fwd")) return rc;
  B200MOE_REQUIRE(context_dim > 0 && n_fields * context_dim <= 8192 && B > 0 && E >= 1 && E <= 64, B200MOE_EINVAL,
                  "ctx_bias_fwd: bad shape (context_dim=%d, B=%d, E=%d)", context_dim, B, E);
  if (int rc = b200moe_check_device()) return rc;
  const int threads = E <= 4 ? 128 : (E <= 8 ? 256 : 512);      // a warp per expert, up to 16 warps
  ctx_bias_fwd_kernel<<<B, threads, sizeof(float) * n_fields * context_dim, static_cast<cudaStream_t>(stream)>>>(
      context, tb, n_fields, context_dim, wc, E, ctx_bias);
  return check_launch("ctx_bias_fwd");
}

extern "C" int b200moe_ctx_bias_bwd(const int64_t* context, const float* dctx_bias, const float* const* tables,
                                    const int32_t* vocab, int n_fields, int context_dim, const float* wc, int B,
                                    int E, float* const* dtables, float* dwc, void* stream) {
  CtxTables tb;
  if (int rc = fill_tables(&tb, tables, dtables, vocab, n_fields, "ctx_bias_bwd")) return rc;
  B200MOE_REQUIRE(dtables != nullptr && dwc != nullptr && dctx_bias != nullptr, B200MOE_EINVAL, "ctx_bias_bwd: NULL output");
  for (int i = 0; i < n_fields; ++i)
    B200MOE_REQUIRE(dtables[i] != nullptr, B200MOE_EINVAL, "ctx_bias_bwd: dtables[%d] is NULL", i);
  if (int rc = b200moe_check_device()) return rc;
  int n_tab = 0;
  for (int i = 0; i < n_fields; ++i) n_tab += vocab[i] * context_dim;
  const int n_dwc = E * n_fields * context_dim;
  const int total = n_dwc + n_tab;
  ctx_bias_bwd_kernel<<<(total + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      context, dctx_bias, tb, n_fields, context_dim, wc, B, E, dwc, n_dwc, n_tab);
  return check_launch("ctx_bias_bwd");
}

extern "C" int b200moe_layer_pdl(int enable) {
  g_pdl_enabled = enable != 0;
  return B200MOE_OK;
}

extern "C" int b200moe_layer_fork(int mode) {
  B200MOE_REQUIRE(mode == 0 || mode == 1, B200MOE_EINVAL, "layer_fork: mode=%d must be 0 or 1", mode);
  g_fork_mode = mode;
  return B200MOE_OK;
}

extern "C" int b200moe_layer_profile(int enable) {
  for (int i = 0; i < g_prof_n; ++i) { cudaEventDestroy(g_prof[i].a); cudaEventDestroy(g_prof[i].b); }
  g_prof_n = 0;
  g_prof_on = enable != 0;
  return B200MOE_OK;
}

extern "C" int b200moe_layer_profile_read(int index, const char** name, float* ms) {
  if (index < 0 || index >= g_prof_n || !name || !ms) return B200MOE_EINVAL;
  *name = g_prof[index].name;
  cudaError_t e = cudaEventElapsedTime(ms, g_prof[index].a, g_prof[index].b);
  if (e != cudaSuccess) { set_error("layer_profile_read: %s", cudaGetErrorString(e)); cudaGetLastError(); return B200MOE_ECUDA; }
  return B200MOE_OK;
}

extern "C" int64_t b200moe_layer_ws_layout(const b200moe_layer_desc* q, int64_t* off) {
  if (check_desc(q, "layer_ws_layout") != B200MOE_OK || off == nullptr) return -1;
  const int64_t T = q->T, d = q->d, F = q->F, E = q->E, k = q->k, R = q->rows_cap;
  const int64_t nblk = b200moe_num_router_blocks(q->T);
  int64_t bytes[B200MOE_WS_COUNT];
  bytes[B200MOE_WS_LOGITS] = T * E * 4;
  bytes[B200MOE_WS_TOPK_IDX] = T * k * 4;
  bytes[B200MOE_WS_TOPK_GATE] = T * k * 4;
  bytes[B200MOE_WS_BLOCK_COUNTS] = nblk * k * E * 4;
  bytes[B200MOE_WS_BLOCK_STATS] = nblk * (E + 1) * 4;
  bytes[B200MOE_WS_EXPERT_OFFSETS] = (E + 1) * 4;
  bytes[B200MOE_WS_META] = B200MOE_META_INTS * 4;
  bytes[B200MOE_WS_TILE_EXPERT] = (R / kRowPad) * 4;
  bytes[B200MOE_WS_TOKENS_PER_EXPERT] = E * 4;
  bytes[B200MOE_WS_KEPT_PER_EXPERT] = E * 4;
  bytes[B200MOE_WS_LOSSES] = 3 * 4;
  bytes[B200MOE_WS_SLOT_LIMIT] = 2 * k * E * 4;
  bytes[B200MOE_WS_ROW_OF_ASSIGNMENT] = T * k * 4;
  bytes[B200MOE_WS_TOKEN_OF_ROW] = R * 4;
  bytes[B200MOE_WS_GATE_OF_ROW] = R * 4;
  bytes[B200MOE_WS_XP] = gather_dispatch(q) ? 0 : R * d * 2;
  bytes[B200MOE_WS_PRE] = R * F * 2;
  bytes[B200MOE_WS_H] = R * F * 2;
  bytes[B200MOE_WS_YO] = R * d * 2;
  bytes[B200MOE_WS_Y_ACC] = reduce_combine(q) ? T * d * 4 : 0;
  bytes[B200MOE_WS_GRID_SYNC] = fused_combine(q) ? 16 : 0;
  bytes[B200MOE_WS_CTX_BIAS] = ((T + q->tokens_per_image - 1) / (q->tokens_per_image > 0 ? q->tokens_per_image : T)) * E * 4;
  int64_t at = 0;
  for (int i = 0; i < B200MOE_WS_COUNT; ++i) { off[i] = at; at += align256(bytes[i]); }
  return at;
}

extern "C" int64_t b200moe_layer_bwd_ws_layout(const b200moe_layer_desc* q, int64_t* off) {
  if (check_desc(q, "layer_bwd_ws_layout") != B200MOE_OK || off == nullptr) return -1;
  const int64_t T = q->T, d = q->d, F = q->F, E = q->E, R = q->rows_cap;
  int64_t bytes[B200MOE_BWS_COUNT];
  bytes[B200MOE_BWS_DYP] = R * d * 2;
  bytes[B200MOE_BWS_DGATE_ROW] = R * 4;
  bytes[B200MOE_BWS_DHPRE] = R * F * 2;
  bytes[B200MOE_BWS_DXP] = R * d * 2;
  bytes[B200MOE_BWS_DLOGITS] = T * E * 4;
  bytes[B200MOE_BWS_DWG_SCRATCH] = b200moe_dwg_scratch_floats(q->T, q->d, q->E) * 4;
  bytes[B200MOE_BWS_DCTX] = ((T + q->tokens_per_image - 1) / (q->tokens_per_image > 0 ? q->tokens_per_image : T)) * E * 4;
  int64_t at = 0;
  for (int i = 0; i < B200MOE_BWS_COUNT; ++i) { off[i] = at; at += align256(bytes[i]); }
  return at;
}

#define WS(T_, i) reinterpret_cast<T_*>(base + off[i])

extern "C" int b200moe_layer_fwd(const b200moe_layer_desc* q, const void* x, const float* ctx_bias,
                                 const b200moe_ctx_desc* cx, const void* wg_split, const void* w1, const float* b1, const void* w2,
                                 const float* b2, void* ws, void* y, int save_for_backward,
                                 const b200moe_postnorm* pn, const b200moe_weight_prep* prep, void* stream) {
  if (int rc = check_desc(q, "layer_fwd")) return rc;
  B200MOE_REQUIRE(x && wg_split && w1 && b1 && w2 && b2 && ws && y, B200MOE_EINVAL, "layer_fwd: NULL argument");
  int64_t off[B200MOE_WS_COUNT];
  b200moe_layer_ws_layout(q, off);
  unsigned char* base = static_cast<unsigned char*>(ws);
  const int T = q->T, d = q->d, F = q->F, E = q->E, k = q->k, R = q->rows_cap;
  const int32_t* meta = WS(int32_t, B200MOE_WS_META);
  const int32_t* tile_expert = WS(int32_t, B200MOE_WS_TILE_EXPERT);
  const int pad = b200moe_layer_row_pad(T, k, E, q->capacity, d, F);
  int rc;
  PdlScope pdl_;
  B200MOE_REQUIRE(!(ctx_bias && cx), B200MOE_EINVAL, "layer_fwd: pass ctx_bias or ctx, not both");
  // weight casts on the caller's side stream: forked here (no dependency on the router), LAUNCHED after the router
  cudaEvent_t weights_ready = prep ? static_cast<cudaEvent_t>(prep->ready) : nullptr;
  const bool cast_here = prep && (prep->w1_f32 || prep->w2_f32);
  ForkState* pf = nullptr;
  if (cast_here) {
    B200MOE_REQUIRE(prep->side_stream != nullptr && prep->side_stream != stream, B200MOE_EINVAL,
                    "layer_fwd: the weight preparation needs a side stream");
    pf = fork_state();
    B200MOE_REQUIRE(pf != nullptr, B200MOE_ECUDA, "layer_fwd: cannot create the fork events");
    cudaEventRecord(pf->ev[0], static_cast<cudaStream_t>(stream));
    cudaStreamWaitEvent(static_cast<cudaStream_t>(prep->side_stream), pf->ev[0], 0);
  }
  auto launch_cast = [&]() -> int {
    if (!cast_here) return B200MOE_OK;
    const bool pdl_was = g_pdl;
    g_pdl = false;
    int rc_ = b200moe_prepare_weights(prep->w1_f32, const_cast<void*>(w1), prep->w1_f32 ? prep->n1 : 0, prep->w2_f32,
                                      const_cast<void*>(w2), prep->w2_f32 ? prep->n2 : 0, nullptr, nullptr, q->d, q->E,
                                      prep->side_stream);
    g_pdl = pdl_was;
    if (rc_) return rc_;
    cudaEventRecord(pf->ev[1], static_cast<cudaStream_t>(prep->side_stream));
    weights_ready = pf->ev[1];
    return B200MOE_OK;
  };
  if (cx && (q->tokens_per_image < 64 || cx->n_fields * cx->context_dim > 256)) {
    // few tokens per image (a 128-token tile would touch many images) or a wide context embedding: the stand-alone
    // kernel fills the workspace slot
    STEP("ctx_bias_fwd", b200moe_ctx_bias_fwd(cx->context, cx->tables, cx->vocab, cx->n_fields, cx->context_dim, cx->wc,
                                              cx->B, E, WS(float, B200MOE_WS_CTX_BIAS), stream));
    ctx_bias = WS(float, B200MOE_WS_CTX_BIAS);
    cx = nullptr;
  }
  if (cx) {
    STEP("router_fwd",
         b200moe_router_fwd_ctx(x, cx, q->tokens_per_image, const_cast<void*>(wg_split), T, d, E, k, q->normalize_topk,
                                WS(float, B200MOE_WS_LOGITS), WS(int32_t, B200MOE_WS_TOPK_IDX),
                                WS(float, B200MOE_WS_TOPK_GATE), WS(int32_t, B200MOE_WS_BLOCK_COUNTS),
                                WS(float, B200MOE_WS_BLOCK_STATS), stream));
  } else {
    STEP("router_fwd",
         b200moe_router_fwd(x, ctx_bias, q->tokens_per_image, nullptr, const_cast<void*>(wg_split), T, d, E, k,
                            q->normalize_topk, WS(float, B200MOE_WS_LOGITS), WS(int32_t, B200MOE_WS_TOPK_IDX),
                            WS(float, B200MOE_WS_TOPK_GATE), WS(int32_t, B200MOE_WS_BLOCK_COUNTS),
                            WS(float, B200MOE_WS_BLOCK_STATS), stream));
  }
  if (int rc_ = launch_cast()) return rc_;
  STEP("plan",
       b200moe_plan(WS(int32_t, B200MOE_WS_TOPK_IDX), WS(float, B200MOE_WS_TOPK_GATE),
                    WS(int32_t, B200MOE_WS_BLOCK_COUNTS), WS(float, B200MOE_WS_BLOCK_STATS), T, k, E, q->capacity,
                    nullptr, nullptr, pad, R, q->aux_loss_weight, q->z_loss_weight,
                    WS(int32_t, B200MOE_WS_EXPERT_OFFSETS), WS(int32_t, B200MOE_WS_META),
                    WS(int32_t, B200MOE_WS_TILE_EXPERT), WS(int32_t, B200MOE_WS_TOKENS_PER_EXPERT),
                    WS(int32_t, B200MOE_WS_KEPT_PER_EXPERT), WS(float, B200MOE_WS_LOSSES),
                    WS(int32_t, B200MOE_WS_SLOT_LIMIT), WS(int32_t, B200MOE_WS_ROW_OF_ASSIGNMENT),
                    WS(int32_t, B200MOE_WS_TOKEN_OF_ROW), WS(float, B200MOE_WS_GATE_OF_ROW), stream));
  const bool fused = fused_combine(q);
  const bool gather = gather_dispatch(q);
  if (gather) {
    // no dispatch kernel: the up-projection gathers its A rows through token_of_row by TMA (and clears the rendezvous
    // counter of the fused combine on its way)
  } else if (fused) {
    // the dispatch kernel also zeroes the rendezvous counter (and the fp32 accumulator of the reduction mode) of the
    // fused combine: no memset launch
    STEP("permute", b200moe_permute_init(x, WS(int32_t, B200MOE_WS_TOKEN_OF_ROW), meta, R, d, WS(void, B200MOE_WS_XP),
                                         reduce_combine(q) ? WS(float, B200MOE_WS_Y_ACC) : nullptr,
                                         WS(int32_t, B200MOE_WS_GRID_SYNC), T, stream));
  } else {
    STEP("permute", b200moe_permute(x, WS(int32_t, B200MOE_WS_TOKEN_OF_ROW), meta, R, d, WS(void, B200MOE_WS_XP), stream));
  }
  {
    // the bf16 expert weights may still be in preparation on another stream (the router, the plan and the permute do
    // not need them): join here, just before their first reader
    if (weights_ready)
      cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), static_cast<cudaEvent_t>(weights_ready), 0);
    PdlOffIf np_(weights_ready != nullptr);
    if (gather) {
      STEP("ffn_up", b200moe_ffn_up_gather(x, WS(int32_t, B200MOE_WS_TOKEN_OF_ROW), T, w1, b1, meta, tile_expert, R, d,
                                           F, E, save_for_backward ? WS(void, B200MOE_WS_PRE) : nullptr,
                                           WS(void, B200MOE_WS_H), q->activation, pad,
                                           fused ? WS(int32_t, B200MOE_WS_GRID_SYNC) : nullptr, stream));
    } else {
      STEP("ffn_up", b200moe_ffn_up(WS(void, B200MOE_WS_XP), w1, b1, meta, tile_expert, R, d, F, E,
                                    save_for_backward ? WS(void, B200MOE_WS_PRE) : nullptr, WS(void, B200MOE_WS_H),
                                    q->activation, pad, stream));
    }
  }
  if (fused) {
    // y -- or, with the fused post-norm, LayerNorm(residual + y) -- leaves the kernel's own tail: the forward ends here.
    // Reduction mode stores yo only for the backward (dgate = <dy, yo>).
    const bool red = reduce_combine(q);
    STEP("ffn_down", b200moe_ffn_down_combine(WS(void, B200MOE_WS_H), w2, b2, meta, tile_expert,
                                              WS(int32_t, B200MOE_WS_TOKEN_OF_ROW), WS(float, B200MOE_WS_GATE_OF_ROW),
                                              WS(int32_t, B200MOE_WS_ROW_OF_ASSIGNMENT), WS(float, B200MOE_WS_TOPK_GATE),
                                              T, k, R, d, F, E,
                                              (!red || save_for_backward) ? WS(void, B200MOE_WS_YO) : nullptr,
                                              red ? WS(float, B200MOE_WS_Y_ACC) : nullptr,
                                              WS(int32_t, B200MOE_WS_GRID_SYNC), y, pn, pad, stream));
    return B200MOE_OK;
  }
  STEP("ffn_down", b200moe_ffn_down(WS(void, B200MOE_WS_H), w2, b2, meta, tile_expert, R, d, F, E,
                                    WS(void, B200MOE_WS_YO), pad, stream));
  if (pn) {
    STEP("combine_ln", b200moe_combine_ln_fwd(WS(void, B200MOE_WS_YO), WS(int32_t, B200MOE_WS_ROW_OF_ASSIGNMENT),
                                              WS(float, B200MOE_WS_TOPK_GATE), nullptr, pn->residual, pn->residual_f32,
                                              pn->gamma, pn->beta, pn->eps, T, k, d, y, pn->out_f32, pn->mean_rstd,
                                              stream));
  } else {
    STEP("combine", b200moe_combine(WS(void, B200MOE_WS_YO), WS(int32_t, B200MOE_WS_ROW_OF_ASSIGNMENT),
                                    WS(float, B200MOE_WS_TOPK_GATE), T, k, d, y, stream));
  }
  return B200MOE_OK;
}

extern "C" int b200moe_layer_bwd(const b200moe_layer_desc* q, const void* x, const float* wg, const void* w1,
                                 const void* w2, const void* ws, void* bws, const void* dy, const float* d_aux,
                                 void* dx, float* dctx_bias, float* dwg, float* dw1, float* db1, float* dw2,
                                 float* db2, int has_ctx_bias, const b200moe_ctx_desc* cx,
                                 const b200moe_postnorm_bwd* pn, void* stream) {
  if (int rc = check_desc(q, "layer_bwd")) return rc;
  B200MOE_REQUIRE(x && wg && w1 && w2 && ws && bws && dy, B200MOE_EINVAL, "layer_bwd: NULL argument");
  int64_t off[B200MOE_WS_COUNT], boff[B200MOE_BWS_COUNT];
  b200moe_layer_ws_layout(q, off);
  b200moe_layer_bwd_ws_layout(q, boff);
  unsigned char* base = static_cast<unsigned char*>(const_cast<void*>(ws));
  unsigned char* bbase = static_cast<unsigned char*>(bws);
#define BWS(T_, i) reinterpret_cast<T_*>(bbase + boff[i])
  const int T = q->T, d = q->d, F = q->F, E = q->E, k = q->k, R = q->rows_cap;
  const int32_t* meta = WS(int32_t, B200MOE_WS_META);
  const int32_t* tile_expert = WS(int32_t, B200MOE_WS_TILE_EXPERT);
  const int32_t* offs = WS(int32_t, B200MOE_WS_EXPERT_OFFSETS);
  const int32_t* roa = WS(int32_t, B200MOE_WS_ROW_OF_ASSIGNMENT);
  void* dyp = BWS(void, B200MOE_BWS_DYP);
  void* dhpre = BWS(void, B200MOE_BWS_DHPRE);
  float* dlogits = BWS(float, B200MOE_BWS_DLOGITS);
  const int pad = b200moe_layer_row_pad(T, k, E, q->capacity, d, F);
  int rc;
  PdlScope pdl_;
  if (pn) {
    STEP("combine_ln_bwd",
         b200moe_combine_ln_bwd(dy, pn->dout_f32, WS(void, B200MOE_WS_YO), roa, WS(float, B200MOE_WS_TOPK_GATE),
                                reduce_combine(q) ? WS(float, B200MOE_WS_Y_ACC) : nullptr, pn->residual,
                                pn->residual_f32, pn->gamma, pn->mean_rstd, T, k, d, pn->dv, pn->dres_f32, pn->scratch,
                                pn->dgamma, pn->dbeta, stream));
    dy = pn->dv;
  }
  ForkState* fk = (!g_prof_on && g_fork_mode) ? fork_state() : nullptr;
  cudaStream_t main_s = static_cast<cudaStream_t>(stream);
  void* s0 = fk ? static_cast<void*>(fk->s0) : stream;   // dlogits alone, high priority
  void* s1 = fk ? static_cast<void*>(fk->s1) : stream;   // router parameter gradients (+ context-path gradients)
  void* s2 = fk ? static_cast<void*>(fk->s2) : stream;   // weight gradients
  void* s3 = fk ? static_cast<void*>(fk->s3) : stream;   // bias column sums
  float* dctx = (dctx_bias && has_ctx_bias) ? dctx_bias : nullptr;
  if (cx != nullptr) {                               // parameter gradients of the context path wanted: dctx is scratch
    B200MOE_REQUIRE(cx->dwc != nullptr && cx->context != nullptr, B200MOE_EINVAL, "layer_bwd: bad context descriptor");
    if (dctx == nullptr) dctx = BWS(float, B200MOE_BWS_DCTX);
  }
  STEP("combine_bwd",
       b200moe_combine_bwd(dy, WS(void, B200MOE_WS_YO), WS(int32_t, B200MOE_WS_TOKEN_OF_ROW),
                           WS(float, B200MOE_WS_GATE_OF_ROW), meta, R, d, dyp, BWS(float, B200MOE_BWS_DGATE_ROW), stream));
  if (fk) {
    cudaEventRecord(fk->ev[0], main_s);
    cudaStreamWaitEvent(fk->s0, fk->ev[0], 0);
    cudaStreamWaitEvent(fk->s2, fk->ev[0], 0);
  }
  // dlogits is queued BEFORE dgrad_down: in a replayed graph a side-branch kernel queued behind a persistent GEMM started
  // only when that GEMM drained (round-2 timeline at 3200 tokens: ready at 58 us, started at 70 us)
  {
    // (tried in round 2 and dropped: finishing dlogits inside combine_bwd by per-token arrival counters -- it removes this
    // launch and the cross-stream edge into unpermute_bwd, but the fence / atomic / dependent-load chain made combine_bwd,
    // which IS on the dx chain, 3.9 -> 12.7 us at 3200 tokens: 104 -> 111 us per step)
    PdlOffIf np_(fk != nullptr);
    STEP("router_bwd_dlogits",
         b200moe_router_bwd_dlogits(WS(float, B200MOE_WS_LOGITS), WS(int32_t, B200MOE_WS_TOPK_IDX), roa,
                                    BWS(float, B200MOE_BWS_DGATE_ROW), WS(int32_t, B200MOE_WS_TOKENS_PER_EXPERT), d_aux,
                                    T, E, k, q->normalize_topk, q->aux_loss_weight, q->z_loss_weight, dlogits, s0));
  }
  if (fk) {
    cudaEventRecord(fk->ev[2], fk->s0);
    cudaStreamWaitEvent(fk->s1, fk->ev[2], 0);     // the rest of the router branch; also joins s0 into s1
  }
  STEP("ffn_dgrad_down",
       b200moe_ffn_dgrad_down(dyp, w2, WS(void, B200MOE_WS_PRE), meta, tile_expert, R, d, F, E, dhpre, q->activation,
                              pad, stream));
  if (fk) cudaEventRecord(fk->ev[1], main_s);
  if (dwg || dctx) {
    STEP("router_bwd_params", b200moe_router_bwd_params(dlogits, x, T, d, E, q->tokens_per_image,
                                                        BWS(float, B200MOE_BWS_DWG_SCRATCH), dwg, dctx, s1));
  }
  if (cx != nullptr) {
    // d(ctx_embed), d(ctx_proj) from dctx, on the router branch (off the dx chain)
    STEP("ctx_bias_bwd", b200moe_ctx_bias_bwd(cx->context, dctx, cx->tables, cx->vocab, cx->n_fields, cx->context_dim,
                                              cx->wc, cx->B, E, cx->dtables, cx->dwc, s1));
  }
  // bias column sums on a branch of their own, as soon as dgrad_down has written dhpre: behind the router gradients
  // (round 1) they were the tail of the whole pass at 3200 tokens (graph timeline: dx done at 96 us, colsum at 107 us),
  // behind the two weight-gradient GEMMs they would be at large rows-per-expert
  if (fk) cudaStreamWaitEvent(fk->s3, fk->ev[1], 0);      // unconditional: the join below needs s3 inside the capture
  {
    PdlOffIf np_(fk != nullptr);
    if (db2 && db1) {
      STEP("colsum", b200moe_colsum(dyp, d, db2, dhpre, F, db1, offs, E, 0, s3));
    } else if (db2) {
      STEP("colsum", b200moe_colsum(dyp, d, db2, nullptr, 0, nullptr, offs, E, 0, s3));
    } else if (db1) {
      STEP("colsum", b200moe_colsum(dhpre, F, db1, nullptr, 0, nullptr, offs, E, 0, s3));
    }
  }
  if (dw2 && dw1 && !gather_dispatch(q)) {
    // both weight gradients in one persistent grid, once dgrad_down has written dhpre
    if (fk) cudaStreamWaitEvent(fk->s2, fk->ev[1], 0);
    PdlOffIf np_(fk != nullptr);
    STEP("ffn_wgrad", b200moe_ffn_wgrad2(dyp, WS(void, B200MOE_WS_H), d, F, dw2, dhpre, WS(void, B200MOE_WS_XP), F, d, dw1,
                                         offs, R, E, 0, s2));
  } else {
    if (dw2) {
      PdlOffIf np_(fk != nullptr);
      STEP("ffn_wgrad", b200moe_ffn_wgrad(dyp, WS(void, B200MOE_WS_H), offs, R, d, F, E, dw2, 0, s2));
    }
    if (fk) cudaStreamWaitEvent(fk->s2, fk->ev[1], 0);
    if (dw1) {
      PdlOffIf np_(fk != nullptr);
      if (gather_dispatch(q)) {
        STEP("ffn_wgrad", b200moe_ffn_wgrad_gather(dhpre, x, WS(int32_t, B200MOE_WS_TOKEN_OF_ROW), T, offs, R, F, d, E,
                                                   dw1, 0, s2));
      } else {
        STEP("ffn_wgrad", b200moe_ffn_wgrad(dhpre, WS(void, B200MOE_WS_XP), offs, R, F, d, E, dw1, 0, s2));
      }
    }
  }
  if (dx) {
    void* dxp = BWS(void, B200MOE_BWS_DXP);
    STEP("ffn_dgrad_up", b200moe_ffn_dgrad_up(dhpre, w1, meta, tile_expert, R, d, F, E, dxp, pad, stream));
    if (fk) cudaStreamWaitEvent(main_s, fk->ev[2], 0);
    PdlOffIf np_(fk != nullptr);
    STEP("unpermute_bwd", b200moe_unpermute_bwd(dxp, roa, dlogits, wg, T, k, d, E, dx, stream));
  }
  if (fk) {
    cudaEventRecord(fk->ev[3], fk->s1);
    cudaEventRecord(fk->ev[4], fk->s2);
    cudaEventRecord(fk->ev[5], fk->s3);
    cudaStreamWaitEvent(main_s, fk->ev[3], 0);
    cudaStreamWaitEvent(main_s, fk->ev[4], 0);
    cudaStreamWaitEvent(main_s, fk->ev[5], 0);
  }
  return B200MOE_OK;
#undef BWS
}
