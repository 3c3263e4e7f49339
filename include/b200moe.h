/* libb200moe — C ABI of the B200-native context-gated MoE FFN layer (sm_100a only).
 *
 * The reference repository (scaleoutsystems/multimodal-MoE) has no FFI for this path: the op this library
 * replaces is the dense encoder FFN `linear2(dropout(gelu(linear1(src))))` at
 *   third_party/rtdetr/rtdetrv2_pytorch/src/zoo/rtdetr/hybrid_encoder.py:129-131 (parameters), :157 (forward)
 * and the layer semantics are SURVEY.md Appendix A (A.1..A.12).  Each entry point below names the step of
 * that path it implements.  A maintainer binds these with ctypes; the Python host class
 * `b200moe.ContextMoE` is the drop-in for hybrid_encoder.py:157.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller (PyTorch caching allocator); the library allocates
 *     nothing and keeps no state between calls;
 *   - every call enqueues work on `stream` (a cudaStream_t passed as void*) and never synchronises;
 *   - return value: 0 = ok, negative = B200MOE_E* code; b200moe_last_error() gives the message (thread-local);
 *   - bf16 tensors are passed as `const void*` / `void*` (row-major, 16-byte aligned rows);
 *   - T tokens, d model width, F expert hidden width, E experts, k = top_k, R = rows of the permuted buffer.
 *   - there is NO CPU fallback: on a device that is not compute capability 10.x every compute call fails.
 */
#ifndef B200MOE_H_
#define B200MOE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200MOE_OK 0
#define B200MOE_EINVAL (-1)   /* bad shape / alignment / argument */
#define B200MOE_EARCH (-2)    /* not an sm_100 device */
#define B200MOE_ECUDA (-3)    /* CUDA runtime / driver error */

/* Expert activation: exact-erf GELU of the encoder FFN (hybrid_encoder.py:157 with zoo/rtdetr/utils.py:160-161) or the
 * ReLU of the decoder FFN (rtdetrv2_decoder.py:169,187,203: activation="relu"). */
#define B200MOE_ACT_GELU 0
#define B200MOE_ACT_RELU 1

#define B200MOE_ROW_PAD 128          /* expert segments of the permuted buffer are padded to this many rows */
#define B200MOE_ROUTER_BLOCK 128     /* max tokens per router block (granularity of the priority scan) */
#define B200MOE_META_INTS 8          /* meta[0]=R (padded rows in use) meta[1]=R/128 meta[2]=dropped meta[3]=kept */

int b200moe_version(void);
/* Kernels launched by the library since load (or since the last call with reset != 0): the evidence behind a
 * benchmark's launch count. */
int64_t b200moe_launch_count(int reset);
const char* b200moe_last_error(void);
/* 0 when the current device is compute capability 10.x, else B200MOE_EARCH. */
int b200moe_check_device(void);

/* Upper bound (multiple of 128) on rows of the permuted buffer: min(T*k, E*capacity) + E*127, rounded up. */
int b200moe_rows_upper_bound(int T, int k, int E, int capacity);
/* Row padding of the expert segments the whole-layer calls use for this problem (128, or 256 when the GEMMs are large
 * enough for CTA pairs) and the matching bound on the rows of the permuted buffer (multiple of row_pad). */
int b200moe_layer_row_pad(int T, int k, int E, int capacity, int d, int F);
int b200moe_layer_rows_cap(int T, int k, int E, int capacity, int d, int F);
/* Tokens per router block (128 = one tcgen05 M tile of the router) and the resulting block count
 * ceil(T / block): first dimension of block_counts / block_stats. */
int b200moe_router_block(int T);
int b200moe_num_router_blocks(int T);

/* fp32 master weights -> bf16 operands (autocast step that torch does before hybrid_encoder.py:157). */
int b200moe_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream);
/* Once per optimizer step, ONE launch: w1, w2 fp32 -> bf16 operands (n1, n2 elements, multiples of 8; either may be
 * 0) and wg [E,d] fp32 -> wg_split (bf16 hi/lo rows for the tcgen05 router, b200moe_router_scratch_bytes; wg may be
 * NULL).  b200moe_router_fwd called with wg == NULL then uses wg_split as it is. */
int b200moe_prepare_weights(const float* w1, void* w1_bf16, int64_t n1, const float* w2, void* w2_bf16, int64_t n2,
                            const float* wg, void* wg_split, int d, int E, void* stream);

/* A.2-A.4 + statistics for A.8/A.9.  Gate GEMV over the token plus the per-image context bias, fp32 softmax,
 * top-k with (p desc, index asc) order, optional renormalisation.
 *   x [T,d] bf16 (d % 64 == 0); ctx_bias [B,E] fp32 or NULL (tok2img = t / tokens_per_image); wg [E,d] fp32
 *   wg_split: scratch of b200moe_router_scratch_bytes(d,E) bytes (the gate product runs on tcgen05 with wg split
 *   into bf16 hi + lo, fp32 accumulate)
 *   logits [T,E] fp32; topk_idx [T,k] int32; topk_gate [T,k] fp32
 *   block_counts [nblk,k,E] int32 (pre-drop assignments per router block, slot, expert)
 *   block_stats  [nblk,E+1] fp32 (sum_t p[t,e] per block; last column sum_t lse^2)                         */
int64_t b200moe_router_scratch_bytes(int d, int E);   /* size of wg_split: bf16 hi/lo split of wg, [2*EP, d] */
/* The context of a batch and the parameters of the A.2 context path, for the calls that evaluate the bias
 * ctx_bias[b,e] = sum_i Embed_i[context[b,i],:] . Wc[e, i*cd:(i+1)*cd] themselves: the router computes the bias of the
 * (at most four) images a 128-token tile touches while its first token tile is in flight, the whole-layer backward
 * appends the parameter gradients to its router branch -- no separate launch either way.  A HOST struct of device
 * pointers. */
#define B200MOE_MAX_CTX_FIELDS 8
typedef struct b200moe_ctx_desc {
  const int64_t* context;                           /* [B, n_fields] ids, one per field per image */
  const float* tables[B200MOE_MAX_CTX_FIELDS];      /* table i: [vocab[i], context_dim] fp32 */
  int32_t vocab[B200MOE_MAX_CTX_FIELDS];
  int32_t n_fields, context_dim, B;
  const float* wc;                                  /* [E, n_fields*context_dim] fp32 */
  float* dtables[B200MOE_MAX_CTX_FIELDS];           /* backward outputs, same shapes (unused by the forward) */
  float* dwc;
} b200moe_ctx_desc;
/* ctx: NULL, or (with ctx_bias == NULL) the bias is evaluated inside the kernel; needs tokens_per_image >= 64. */
int b200moe_router_fwd_ctx(const void* x, const b200moe_ctx_desc* ctx, int tokens_per_image, void* wg_split, int T,
                           int d, int E, int k, int normalize_topk, float* logits, int32_t* topk_idx,
                           float* topk_gate, int32_t* block_counts, float* block_stats, void* stream);
int b200moe_router_fwd(const void* x, const float* ctx_bias, int tokens_per_image, const float* wg,
                       void* wg_split, int T, int d, int E, int k, int normalize_topk,
                       float* logits, int32_t* topk_idx, float* topk_gate,
                       int32_t* block_counts, float* block_stats, void* stream);

/* A.5 + A.8-A.10.  Slot-major / token-major priority scan, capacity clamp, padded expert offsets, row maps, losses.
 *   global_offset [k,E] int32 or NULL: earlier-priority assignments owned by other ranks (expert parallelism)
 *   expert_order  [E] int32 or NULL : position of expert e in the row layout (NULL = identity)
 *   row_pad: 128 or 256 on one GPU (GEMM tiles / CTA-pair tiles), 1 for an all-to-all send buffer
 *   out: expert_offsets [E+1] (in layout order: offsets[i] is the first row of the expert at position i)
 *        meta [B200MOE_META_INTS]; tile_expert [rows_cap/128] (expert of each 128-row tile, -1 beyond R)
 *        tokens_per_expert [E] (pre-drop); kept_per_expert [E]
 *        losses [3] = {aux_loss_weight*lb + z_loss_weight*z, lb, z}
 *        row_of_assignment [T,k] (-1 = dropped); token_of_row [rows_cap] (-1 = pad); gate_of_row [rows_cap]
 *   block_counts is scratch after the call (large T: overwritten with its exclusive scan over the blocks); slot_limit [2,k,E] is scratch. */
int b200moe_plan(const int32_t* topk_idx, const float* topk_gate, int32_t* block_counts,
                 const float* block_stats, int T, int k, int E, int capacity,
                 const int32_t* global_offset, const int32_t* expert_order, int row_pad, int rows_cap,
                 float aux_loss_weight, float z_loss_weight,
                 int32_t* expert_offsets, int32_t* meta, int32_t* tile_expert,
                 int32_t* tokens_per_expert, int32_t* kept_per_expert, float* losses, int32_t* slot_limit,
                 int32_t* row_of_assignment, int32_t* token_of_row, float* gate_of_row, void* stream);

/* Dispatch: xp[r,:] = x[token_of_row[r],:] (zero rows for pads), r < meta[0]. 128-bit loads/stores. */
int b200moe_permute(const void* x, const int32_t* token_of_row, const int32_t* meta, int rows_cap, int d,
                    void* xp, void* stream);
/* The same dispatch kernel, also preparing the fused combine of b200moe_ffn_down_combine (no memset launches on the
 * critical path): grid_sync (16 bytes, 16-byte aligned) := 0 and, for the reduction mode, y_acc [T,d] fp32 := 0
 * (either may be NULL). */
int b200moe_permute_init(const void* x, const int32_t* token_of_row, const int32_t* meta, int rows_cap, int d,
                         void* xp, float* y_acc, int32_t* grid_sync, int T, void* stream);

/* A.7 combine: y[t,:] = sum_s gate[t,s] * yo[row_of_assignment[t,s],:] in slot order, fp32 accumulate. */
int b200moe_combine(const void* yo, const int32_t* row_of_assignment, const float* topk_gate,
                    int T, int k, int d, void* y, void* stream);

/* Backward of combine: dyp[r,:] = gate_of_row[r] * dy[token_of_row[r],:] (0 for pads);
 * dgate_row[r] = <dy[token,:], yo[r,:]>. */
int b200moe_combine_bwd(const void* dy, const void* yo, const int32_t* token_of_row, const float* gate_of_row,
                        const int32_t* meta, int rows_cap, int d, void* dyp, float* dgate_row, void* stream);

/* Expert FFN, forward (A.6).  Grouped bf16 GEMMs on tcgen05/TMEM fed by TMA.
 *   up  : pre = xp @ w1[e]^T + b1[e];  h = act(pre)           w1 [E,F,d] bf16, b1 [E,F] fp32; act = B200MOE_ACT_*
 *   down: yo  = h  @ w2[e]^T + b2[e]                          w2 [E,d,F] bf16, b2 [E,d] fp32
 *   pre may be NULL (inference: nothing saved for backward).
 *   row_pad = the padding b200moe_plan laid the expert segments out with: 128, or 256 -- then 256-row tiles never
 *   straddle experts and large problems run on CTA pairs (tcgen05 cta_group::2: two SMs share one 256 x 256 tile,
 *   each loading half of the weight tile).  The weight-gradient GEMM pairs along the weight rows by itself.   */
int b200moe_ffn_up(const void* xp, const void* w1, const float* b1, const int32_t* meta,
                   const int32_t* tile_expert, int rows_cap, int d, int F, int E,
                   void* pre, void* h, int activation, int row_pad, void* stream);
/* The up-projection with the dispatch gather of A.5 fused into its A-operand load: row r of the permuted buffer is
 * x[token_of_row[r], :] fetched by the kernel's own TMA loads (tile::gather4; pad rows, token_of_row < 0, arrive as
 * zeros), so xp [R,d] is never written and b200moe_permute is not launched.  x [T,d] bf16.  `clear`: NULL, or one int32
 * the kernel sets to zero (the grid_sync counter of the b200moe_ffn_down_combine that follows).  Bit-identical to
 * b200moe_permute + b200moe_ffn_up; slower (see B200MOE_LAYER_GATHER_DISPATCH).                                    */
int b200moe_ffn_up_gather(const void* x, const int32_t* token_of_row, int T, const void* w1, const float* b1,
                          const int32_t* meta, const int32_t* tile_expert, int rows_cap, int d, int F, int E,
                          void* pre, void* h, int activation, int row_pad, int32_t* clear, void* stream);
int b200moe_ffn_down(const void* h, const void* w2, const float* b2, const int32_t* meta,
                     const int32_t* tile_expert, int rows_cap, int d, int F, int E,
                     void* yo, int row_pad, void* stream);
/* Optional fused post-norm of the encoder layer (hybrid_encoder.py:158-160): with a non-NULL b200moe_postnorm the last
 * kernel of the forward computes out = LayerNorm(residual + combine(yo)) instead of y = combine(yo) (the [T,d]
 * intermediates y and y + residual are never written), and the backward starts from d(out).                   */
typedef struct b200moe_postnorm {
  const void* residual;     /* [T,d] bf16, or fp32 with residual_f32 (the encoder's residual stream under autocast) */
  const float* gamma;       /* [d] LayerNorm weight */
  const float* beta;        /* [d] LayerNorm bias */
  float eps;
  float* mean_rstd;         /* [T,2] fp32, written by the forward, read by the backward */
  int32_t residual_f32;     /* dtype of residual: 0 = bf16, 1 = fp32 */
  int32_t out_f32;          /* dtype of the layer output y: 0 = bf16, 1 = fp32 */
} b200moe_postnorm;
typedef struct b200moe_postnorm_bwd {
  const void* residual;     /* as in the forward */
  const float* gamma;
  const float* mean_rstd;   /* from the forward */
  void* dv;                 /* [T,d] bf16 out: gradient of the pre-norm sum = d(MoE output) (and d(residual) when that is bf16) */
  float* scratch;           /* b200moe_ln_scratch_floats(T, d) floats */
  float* dgamma;            /* [d] out */
  float* dbeta;             /* [d] out */
  int32_t residual_f32;
  int32_t dout_f32;         /* dtype of dy = d(out): 0 = bf16, 1 = fp32 */
  float* dres_f32;          /* [T,d] fp32 out or NULL: the same gradient unrounded, for an fp32 residual */
} b200moe_postnorm_bwd;
/* Down-projection and the gate-weighted combine of A.7 in ONE kernel (the `linear2` half of hybrid_encoder.py:157 and
 * the weighted sum over the selected experts; with `postnorm` also :158-160).  Two modes:
 *   gather mode (y_acc == NULL; any top_k): the epilogues store the expert outputs yo [R,d] as b200moe_ffn_down does;
 *     the kernel ends with a grid-wide rendezvous (the grid is persistent: never more CTAs than are resident at once),
 *     after which its CTAs compute y[t,:] = sum_s topk_gate[t,s] * yo[row_of_assignment[t,s],:] (slot order, fp32) from
 *     the rows just written -- bit-identical to b200moe_ffn_down + b200moe_combine, no second launch.
 *   reduction mode (y_acc != NULL; top_k <= 2): the combine happens IN THE EPILOGUE, straight from the accumulator
 *     registers: y_acc[token_of_row[r], :] += gate_of_row[r] * bf16(h[r] @ w2[e]^T + b2[e]) as fp32 reductions at L2.
 *     Two fp32 terms commute, so the result does not depend on the order in which tiles finish and is bit-identical to
 *     the other paths.  yo may be NULL (inference: the backward is its only reader).  y_acc [T,d] fp32 must be zero
 *     on entry (b200moe_permute_init).  With grid_sync + y the tail converts y_acc to y after the rendezvous; with
 *     both NULL the caller finishes from y_acc (b200moe_combine_ln_fwd).  Measured (round 2): the 2 x A x d fp32
 *     reductions cost more than the separate combine kernel from ~10^4 tokens on (L2 atomic throughput), so the
 *     whole-layer call uses gather mode unless B200MOE_LAYER_COMBINE_REDUCE is set.
 *   grid_sync : one int32 counter, zero on entry (b200moe_permute_init)
 *   postnorm  : NULL, or (d <= 1024) the tail writes y = LayerNorm(residual + combine) instead of the combine. */
int b200moe_ffn_down_combine(const void* h, const void* w2, const float* b2, const int32_t* meta,
                             const int32_t* tile_expert, const int32_t* token_of_row, const float* gate_of_row,
                             const int32_t* row_of_assignment, const float* topk_gate,
                             int T, int k, int rows_cap, int d, int F, int E,
                             void* yo, float* y_acc, int32_t* grid_sync, void* y,
                             const b200moe_postnorm* postnorm, int row_pad, void* stream);

/* Expert FFN, backward.
 *   dgrad_down: dhpre = (dyp @ w2[e]) * act'(pre)             [R,F]
 *   dgrad_up  : dxp   = dhpre @ w1[e]                         [R,d]
 *   wgrad     : dw[e] = a[rows of e]^T @ b[rows of e]         a [R,M] bf16, b [R,N] bf16 -> dw [E,M,N] fp32
 *               (dw2 = dyp^T h, dw1 = dhpre^T xp); accumulate != 0 adds into dw
 *   colsum    : db[e,:] = sum over rows of e of a[r,:]        (db2 from dyp, db1 from dhpre)  */
int b200moe_ffn_dgrad_down(const void* dyp, const void* w2, const void* pre, const int32_t* meta,
                           const int32_t* tile_expert, int rows_cap, int d, int F, int E,
                           void* dhpre, int activation, int row_pad, void* stream);
int b200moe_ffn_dgrad_up(const void* dhpre, const void* w1, const int32_t* meta,
                         const int32_t* tile_expert, int rows_cap, int d, int F, int E,
                         void* dxp, int row_pad, void* stream);
int b200moe_ffn_wgrad(const void* a, const void* b, const int32_t* expert_offsets, int rows_cap,
                      int M, int N, int E, float* dw, int accumulate, void* stream);
/* Both weight gradients of the FFN in one launch: dw_a[e] = a0^T b0 [E,M0,N0] and dw_b[e] = a1^T b1 [E,M1,N1]
 * (dw2 = dyp^T h: M0 = d, N0 = F; dw1 = dhpre^T xp: M1 = F, N1 = d), bit-identical to two b200moe_ffn_wgrad calls.  The
 * tiles of both problems share one persistent grid (no partly filled last wave per problem, no queueing of the second
 * launch behind the first). */
int b200moe_ffn_wgrad2(const void* a0, const void* b0, int M0, int N0, float* dw_a,
                       const void* a1, const void* b1, int M1, int N1, float* dw_b,
                       const int32_t* expert_offsets, int rows_cap, int E, int accumulate, void* stream);
/* wgrad with b = x[token_of_row[rows of e], :] gathered by the TMA loads of the B operand (dw1 = dhpre^T xp without
 * the permuted copy; x [T,N] bf16). */
int b200moe_ffn_wgrad_gather(const void* a, const void* x, const int32_t* token_of_row, int T,
                             const int32_t* expert_offsets, int rows_cap, int M, int N, int E, float* dw,
                             int accumulate, void* stream);
/* Bias gradients of both projections in ONE launch (a1 may be NULL): db0[e,:] = column sums of a0 over the rows of
 * expert e, same for (a1, db1).  Fixed summation order, no scratch.  ncols multiples of 64. */
int b200moe_colsum(const void* a0, int ncols0, float* db0, const void* a1, int ncols1, float* db1,
                   const int32_t* expert_offsets, int E, int accumulate, void* stream);

/* A.12 router gradient.  dlogits from the gate gradient (softmax + renormalisation Jacobian on the selected k),
 * the load-balance term through P_e and the z-loss through lse.  d_aux is the upstream gradient of the weighted
 * auxiliary loss (device scalar, may be NULL = 0).  dgate_row may be NULL (no gradient through the gates).    */
int b200moe_router_bwd_dlogits(const float* logits, const int32_t* topk_idx, const int32_t* row_of_assignment,
                               const float* dgate_row, const int32_t* tokens_per_expert, const float* d_aux,
                               int T, int E, int k, int normalize_topk, float aux_loss_weight,
                               float z_loss_weight, float* dlogits, void* stream);
/* dwg [E,d] = dlogits^T x (fp32, deterministic two-stage reduction; scratch >= b200moe_dwg_scratch_floats). */
int64_t b200moe_dwg_scratch_floats(int T, int d, int E);
int b200moe_router_bwd_dwg(const float* dlogits, const void* x, int T, int d, int E, float* scratch,
                           float* dwg, void* stream);
/* The same gradient for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The scratch
 * is a device array built into the library, B200MOE_DWG_STATIC_MAX_FLOATS floats = 16 MB, held by every GPU that loads
 * the library until the process exits; the call allocates nothing (no cudaMalloc / cudaFree, no allocating library
 * routine) and fails with B200MOE_EINVAL when b200moe_dwg_scratch_floats(T, d, E) exceeds the array (e.g. d = 1024 at
 * 16384 tokens; 3200 x 256 x 8 needs 0.4 M floats).  All calls on a device share the array: calls that overlap on the
 * GPU (different streams) race -- serialise them.  Results are bit-identical to b200moe_router_bwd_dwg. */
#define B200MOE_DWG_STATIC_MAX_FLOATS (4 * 1024 * 1024)
int b200moe_router_bwd_dwg_noalloc(const float* dlogits, const void* x, int T, int d, int E, float* dwg,
                                   void* stream);
/* Both parameter gradients of the router in two launches (either output may be NULL): dwg as above and
 * dctx_bias [B,E] = segment sums of dlogits over the tokens of each image, reduced inside the second launch. */
int b200moe_router_bwd_params(const float* dlogits, const void* x, int T, int d, int E, int tokens_per_image,
                              float* scratch, float* dwg, float* dctx_bias, void* stream);
/* dctx_bias [B,E] = segment sums of dlogits over the tokens of each image. */
int b200moe_router_bwd_dctx(const float* dlogits, int T, int E, int tokens_per_image, float* dctx_bias,
                            void* stream);
/* Backward of permute fused with the router's dx: dx[t,:] = sum_s dxp[row_of_assignment[t,s],:] + dlogits[t,:] @ wg.
 * dxp / dlogits may each be NULL. */
int b200moe_unpermute_bwd(const void* dxp, const int32_t* row_of_assignment, const float* dlogits,
                          const float* wg, int T, int k, int d, int E, void* dx, void* stream);

/* Expert-parallel receive side: rows arrive as n_seg <= 1024 back-to-back segments (source rank major, local
 * expert minor); build the 128-row-padded expert-major layout.
 *   recv_counts [n_seg] int32 rows per segment, seg_expert [n_seg] local expert of each segment
 *   out: expert_offsets [E_local+1] (padded to 128), meta, tile_expert, src_of_row [rows_cap] (-1 = pad)      */
int b200moe_ep_recv_plan(const int32_t* recv_counts, const int32_t* seg_expert, int n_seg, int E_local,
                         int rows_cap, int32_t* expert_offsets, int32_t* meta, int32_t* tile_expert,
                         int32_t* src_of_row, void* stream);
/* Expert-parallel routing tables on the device (no host round trip before the split sizes are needed).
 *   ep_hist        : out [k*E+1] = per-(slot, expert) assignment counts of this rank (sum of block_counts over the
 *                    router blocks), then T: the message of the histogram all-gather.
 *   ep_route_tables: all_hist [W, k*E+1] (gathered); capacity_factor < 0 = no limit.  Outputs: global_offset [k,E] for
 *                    b200moe_plan; kept [W*E+2] = rows every rank sends to every expert, then the global capacity and
 *                    the global token count (copy this to the host for the all-to-all split sizes); recv_counts
 *                    [C][W][experts_per_chunk] = rows this rank receives per (chunk of local experts, source, expert). */
int b200moe_ep_hist(const int32_t* block_counts, int nblk, int k, int E, int T, int32_t* out, void* stream);
int b200moe_ep_route_tables(const int32_t* all_hist, int W, int k, int E, int rank, double capacity_factor,
                            int experts_per_chunk, int32_t* global_offset, int32_t* kept, int32_t* recv_counts,
                            void* stream);
/* Inverse row gather used on the EP return path: out[src_of_row[r],:] = in[r,:] for r < meta[0], src >= 0. */
int b200moe_scatter_rows(const void* in, const int32_t* src_of_row, const int32_t* meta, int rows_cap, int d,
                         void* out, void* stream);

/* ---- Expert parallelism over peer memory (NVLink loads / stores inside the row-movement kernels) ----------------
 * Every rank owns an arena holding row buffers [rows_cap, d] bf16 in the padded expert-major layout (xp, yo, dyp,
 * dxp at byte offsets the host chooses); `peers` is a DEVICE array of W arena base addresses as mapped on this GPU
 * (symmetric memory).  A remote row is named dest = rank * rows_cap + row.  The host places one device-side barrier
 * over the ranks between a producer and its remote consumers; nothing here synchronises.
 *   ep_p2p_tables      : kept [W*E+2] from b200moe_ep_route_tables -> dest_off [E] (first row, on rank e / E_local, of
 *                        this rank's rows for expert e), this rank's expert_offsets [E_local+1] / meta / tile_expert
 *                        (as b200moe_plan writes them); meta[5] = 1 if some rank's rows exceed rows_cap.
 *   ep_p2p_zero_pads   : zero the pad rows of up to two local buffers (a, b may be NULL).
 *   ep_p2p_dispatch    : peers' xp[dest] = x[token_of_row[r]] for every send row r (send layout = b200moe_plan with
 *                        row_pad 1); writes dest_of_row [send_rows_cap] for the calls below.
 *   ep_p2p_combine     : y[t] = sum_s gate[t,s] * peers' yo[dest_of_row[row_of_assignment[t,s]]].
 *   ep_p2p_combine_bwd : peers' dyp[dest] = gate_of_row[r] * dy[token]; dgate_row[r] = <dy[token], peers' yo[dest]>.
 *   ep_p2p_unpermute_bwd: b200moe_unpermute_bwd with the dxp rows loaded from the peers.
 *   ep_p2p_pull_rows   : out[r] = peers' buffer[dest_of_row[r]] for every send row (plain copy, no shared memory: runs
 *                        next to the weight-gradient GEMMs on a side stream; b200moe_unpermute_bwd then reads `out`). */
int b200moe_ep_p2p_tables(const int32_t* kept, int W, int E, int rank, int rows_cap, int row_pad /* 128 | 256 */,
                          int32_t* dest_off, int32_t* expert_offsets, int32_t* meta, int32_t* tile_expert, void* stream);
int b200moe_ep_p2p_zero_pads(const int32_t* kept, const int32_t* expert_offsets, int W, int E, int rank, int d,
                             void* a, void* b, void* stream);
int b200moe_ep_p2p_dispatch(const void* x, const int32_t* token_of_row, const int32_t* send_offsets,
                            const int32_t* send_meta, const int32_t* dest_off, const void* const* peers,
                            int64_t buf_offset_bytes, int rows_cap, int send_rows_cap, int E, int E_local, int d,
                            int32_t* dest_of_row, void* stream);
int b200moe_ep_p2p_combine(const void* const* peers, int64_t buf_offset_bytes, int rows_cap,
                           const int32_t* dest_of_row, const int32_t* row_of_assignment, const float* topk_gate,
                           int T, int k, int d, void* y, void* stream);
int b200moe_ep_p2p_combine_bwd(const void* dy, const void* const* peers, int64_t yo_offset_bytes,
                               int64_t dyp_offset_bytes, int rows_cap, const int32_t* dest_of_row,
                               const int32_t* token_of_row, const float* gate_of_row, const int32_t* send_meta,
                               int send_rows_cap, int d, float* dgate_row, void* stream);
int b200moe_ep_p2p_pull_rows(const void* const* peers, int64_t buf_offset_bytes, int rows_cap,
                             const int32_t* dest_of_row, const int32_t* send_meta, int send_rows_cap, int d, void* out,
                             void* stream);
int b200moe_ep_p2p_unpermute_bwd(const void* const* peers, int64_t buf_offset_bytes, int rows_cap,
                                 const int32_t* dest_of_row, const int32_t* row_of_assignment, const float* dlogits,
                                 const float* wg, int T, int k, int d, int E, void* dx, void* stream);

/* ---- Context bias (A.2): ctx_bias[b,e] = sum_i Embed_i[context[b,i],:] . Wc[e, i*cd:(i+1)*cd] ------------------
 * context [B,n_fields] int64 (device); tables: HOST array of n_fields device pointers, table i is [vocab[i], cd]
 * fp32; vocab: HOST array; wc [E, n_fields*cd] fp32.  Ids outside [0, vocab[i]) contribute nothing.
 * Backward writes every element of dtables[i] and dwc (fixed summation order over the images).                 */
int b200moe_ctx_bias_fwd(const int64_t* context, const float* const* tables, const int32_t* vocab, int n_fields,
                         int context_dim, const float* wc, int B, int E, float* ctx_bias, void* stream);
int b200moe_ctx_bias_bwd(const int64_t* context, const float* dctx_bias, const float* const* tables,
                         const int32_t* vocab, int n_fields, int context_dim, const float* wc, int B, int E,
                         float* const* dtables, float* dwc, void* stream);

/* ---- Whole layer in one call each way (what b200moe.ContextMoE uses) ---------------------------------------------
 * The caller owns two workspaces whose layout b200moe_layer_ws_layout / b200moe_layer_bwd_ws_layout describe
 * (byte offsets of every buffer, 256-byte aligned; return value = total bytes, < 0 on a bad descriptor).  The
 * forward workspace doubles as "saved for backward" and as the telemetry source (logits, topk_idx, ...).      */
typedef struct b200moe_layer_desc {
  int32_t T, d, F, E, k;
  int32_t tokens_per_image;   /* tok2img = t / tokens_per_image (ignored when ctx_bias is NULL) */
  int32_t capacity;           /* A.5: ceil(capacity_factor * T * k / E), or T*k for "no limit" */
  int32_t rows_cap;           /* b200moe_rows_upper_bound(T, k, E, capacity) */
  int32_t normalize_topk;
  float aux_loss_weight, z_loss_weight;
  int32_t activation;         /* B200MOE_ACT_GELU | B200MOE_ACT_RELU */
  int32_t flags;              /* B200MOE_LAYER_* */
} b200moe_layer_desc;
/* flags: where the A.7 combine runs.  Default (no bit): inside the down-projection kernel's tail
 * (b200moe_ffn_down_combine, gather mode) while the expert outputs yo [R,d] stay L2-resident (R*d*2 <= 48 MB), the
 * separate kernels beyond that -- measured, round 2: tail 18.5 vs 22.6 us at 3200 tokens, 0.076 vs 0.083 ms at
 * C/14720, but 0.272 vs 0.235 ms at C/65533, where the rendezvous idles the early CTAs longer than a launch costs.
 *   B200MOE_LAYER_SEPARATE_COMBINE: always b200moe_ffn_down + b200moe_combine (the tested reference);
 *   B200MOE_LAYER_COMBINE_TAIL    : always the tail of the down-projection kernel;
 *   B200MOE_LAYER_COMBINE_REDUCE  : in the epilogue itself by fp32 reductions (top_k <= 2; tail else). */
#define B200MOE_LAYER_SEPARATE_COMBINE 1
#define B200MOE_LAYER_COMBINE_REDUCE 2
#define B200MOE_LAYER_COMBINE_TAIL 4
/* Dispatch.  Default: b200moe_permute writes the permuted tokens xp [R,d] and the GEMMs read them with tiled TMA loads.
 *   B200MOE_LAYER_GATHER_DISPATCH: the permuted copy is never made -- b200moe_ffn_up_gather in the forward and
 *   b200moe_ffn_wgrad_gather in the backward fetch x[token_of_row[r]] through tile::gather4 TMA loads (no
 *   b200moe_permute launch, workspace slot XP empty).  Bit-identical; measured SLOWER on B200 (round 2: a gather4 moves
 *   4 x 128 B and an SM's TMA unit retires one per ~95 cycles, ~5 B/clk against >100 B/clk for tiled boxes: up-projection
 *   0.235 -> 0.79 ms at E=64 d=1024, the w1 weight gradient 0.22 -> 0.78 ms at C/65533), so it is an option, not the
 *   default.  Ignored with COMBINE_REDUCE (whose accumulator the permute kernel clears). */
#define B200MOE_LAYER_GATHER_DISPATCH 8

enum {
  B200MOE_WS_LOGITS = 0, B200MOE_WS_TOPK_IDX, B200MOE_WS_TOPK_GATE, B200MOE_WS_BLOCK_COUNTS, B200MOE_WS_BLOCK_STATS,
  B200MOE_WS_EXPERT_OFFSETS, B200MOE_WS_META, B200MOE_WS_TILE_EXPERT, B200MOE_WS_TOKENS_PER_EXPERT,
  B200MOE_WS_KEPT_PER_EXPERT, B200MOE_WS_LOSSES, B200MOE_WS_SLOT_LIMIT, B200MOE_WS_ROW_OF_ASSIGNMENT,
  B200MOE_WS_TOKEN_OF_ROW, B200MOE_WS_GATE_OF_ROW, B200MOE_WS_XP, B200MOE_WS_PRE, B200MOE_WS_H, B200MOE_WS_YO,
  B200MOE_WS_Y_ACC /* [T,d] fp32 */, B200MOE_WS_GRID_SYNC /* 4 x int32 */, B200MOE_WS_CTX_BIAS /* [B,E] fp32 */,
  B200MOE_WS_COUNT
};
enum {
  B200MOE_BWS_DYP = 0, B200MOE_BWS_DGATE_ROW, B200MOE_BWS_DHPRE, B200MOE_BWS_DXP, B200MOE_BWS_DLOGITS,
  B200MOE_BWS_DWG_SCRATCH, B200MOE_BWS_DCTX /* [B,E] fp32 */, B200MOE_BWS_COUNT
};
/* y_acc != NULL: the combine already happened in the down-projection's epilogue; the pre-norm sum is y_acc + residual
 * (yo / row_of_assignment / topk_gate are then unused and may be NULL). */
int b200moe_combine_ln_fwd(const void* yo, const int32_t* row_of_assignment, const float* topk_gate,
                           const float* y_acc, const void* residual, int residual_f32, const float* gamma,
                           const float* beta, float eps, int T, int k, int d, void* out, int out_f32,
                           float* mean_rstd, void* stream);
int64_t b200moe_ln_scratch_floats(int T, int d);
int b200moe_combine_ln_bwd(const void* dout, int dout_f32, const void* yo, const int32_t* row_of_assignment,
                           const float* topk_gate, const float* y_acc, const void* residual, int residual_f32,
                           const float* gamma, const float* mean_rstd, int T, int k, int d, void* dv, float* dres_f32,
                           float* scratch, float* dgamma, float* dbeta, void* stream);

int64_t b200moe_layer_ws_layout(const b200moe_layer_desc* desc, int64_t* offsets /* [B200MOE_WS_COUNT] */);
int64_t b200moe_layer_bwd_ws_layout(const b200moe_layer_desc* desc, int64_t* offsets /* [B200MOE_BWS_COUNT] */);
/* router -> plan -> permute -> up(+activation) -> down with the combine fused into its epilogue (top_k <= 2; else, or with
 * B200MOE_LAYER_SEPARATE_COMBINE: down -> combine).  x [T,d] bf16; ctx_bias [B,E] fp32 or NULL; wg_split
 * from b200moe_prepare_weights; w1 [E,F,d], w2 [E,d,F] bf16; b1 [E,F], b2 [E,d] fp32; y [T,d] bf16.
 * save_for_backward == 0 skips the pre-activation store. */
/* The context enters either as ctx_bias [B,E] (computed by the caller: b200moe_ctx_bias_fwd) or as ctx (evaluated
 * inside the router kernel; both NULL: no context). */
/* Preparation of the bf16 expert operands overlapped with the first kernels of the pass (which need only the gate
 * split): the fp32 -> bf16 casts of b200moe_prepare_weights, launched by b200moe_layer_fwd itself on `side_stream`
 * AFTER its router kernel -- a cast launched first fills every thread slot of every SM and the router then waits for
 * the cast's last wave (round 2: 204 us at 32 experts of d=1024, F=4096) -- and joined just before the up-projection.
 * w1_f32 / w2_f32 may be NULL individually (that operand is already bf16).  Alternatively `ready` alone: a cudaEvent_t
 * the caller recorded after preparing w1 / w2 on a stream of its own. */
typedef struct b200moe_weight_prep {
  const float* w1_f32; int64_t n1;   /* element counts, multiples of 8; results go to the w1 / w2 passed to the call */
  const float* w2_f32; int64_t n2;
  void* side_stream;                 /* cudaStream_t, not the call's stream */
  void* ready;                       /* cudaEvent_t or NULL */
} b200moe_weight_prep;
int b200moe_layer_fwd(const b200moe_layer_desc* desc, const void* x, const float* ctx_bias,
                      const b200moe_ctx_desc* ctx, const void* wg_split,
                      const void* w1, const float* b1, const void* w2, const float* b2, void* ws, void* y,
                      int save_for_backward, const b200moe_postnorm* postnorm /* or NULL */,
                      const b200moe_weight_prep* prep /* or NULL: w1 / w2 are ready */,
                      void* stream);
/* Mirror of the forward incl. the A.12 router gradient.  dy [T,d] bf16; d_aux: device scalar (gradient of the
 * weighted auxiliary loss) or NULL.  Any of the gradient outputs may be NULL (not computed):
 * dx [T,d] bf16, dctx_bias [B,E], dwg [E,d], dw1 [E,F,d], db1 [E,F], dw2 [E,d,F], db2 [E,d] (all fp32).      */
int b200moe_layer_bwd(const b200moe_layer_desc* desc, const void* x, const float* wg, const void* w1,
                      const void* w2, const void* ws, void* bws, const void* dy, const float* d_aux, void* dx,
                      float* dctx_bias, float* dwg, float* dw1, float* db1, float* dw2, float* db2,
                      int has_ctx_bias, const b200moe_ctx_desc* ctx /* or NULL; with it the gradients ctx->dtables and
                      ctx->dwc are written on the router branch (dctx_bias may then be NULL) */,
                      const b200moe_postnorm_bwd* postnorm /* or NULL: dy is then d(out) */, void* stream);
/* Per-step CUDA-event timing of the two calls above (measurement aid): enable, run, synchronize the stream, then read
 * entries 0,1,... until the call returns B200MOE_EINVAL.  b200moe_layer_profile(x) also clears earlier entries. */
int b200moe_layer_profile(int enable);
/* Programmatic dependent launch between the kernels of a whole-layer call (default on): every kernel of the path runs
 * its prologue while its predecessor drains and executes griddepcontrol.wait before touching global memory. */
int b200moe_layer_pdl(int enable);
/* Backward-pass fork/join: weight gradients + bias sums and the router's parameter gradients run on two library-owned
 * side streams next to the dgrad chain and are joined to `stream` before b200moe_layer_bwd returns (capturable).
 * mode 0 = off, 1 = on (default; measured 3-12 % faster per step from 3200 to 65600 tokens).  The side streams and their
 * events are per device, not per caller: whole-layer calls on one device must come from one host thread at a time. */
int b200moe_layer_fork(int mode);
int b200moe_layer_profile_read(int index, const char** name, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* B200MOE_H_ */
