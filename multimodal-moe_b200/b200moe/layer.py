"""ContextMoE — drop-in for the dense encoder FFN of the vendored RT-DETR layer.

Replaces `linear2(dropout(gelu(linear1(src))))` at
third_party/rtdetr/rtdetrv2_pytorch/src/zoo/rtdetr/hybrid_encoder.py:157 (parameters :129-131) with the
context-gated mixture-of-experts layer of SURVEY.md §8(b) / Appendix A.  Constructor and forward follow §8(b):
`forward(x[..., d_model], context[B, n_fields] | None) -> Tensor[..., d_model]`; the weighted auxiliary loss is left
on `self.aux_loss` (autograd-connected) and telemetry on `self.last_aux`.

All compute is in libb200moe.so (hand-written sm_100a CUDA behind the C ABI of include/b200moe.h).  There is no
PyTorch or CPU fallback: without the library, or on a non-sm_100 device, forward raises.
"""
from __future__ import annotations

import ctypes as C
import math
from collections.abc import Mapping
from typing import Optional, Sequence

import torch
import torch.nn as nn

from . import _lib, ops


_WS_DTYPES = {"logits": torch.float32, "topk_idx": torch.int32, "topk_gate": torch.float32,
              "block_counts": torch.int32, "block_stats": torch.float32, "expert_offsets": torch.int32,
              "meta": torch.int32, "tile_expert": torch.int32, "tokens_per_expert": torch.int32,
              "kept_per_expert": torch.int32, "losses": torch.float32, "slot_limit": torch.int32,
              "row_of_assignment": torch.int32, "token_of_row": torch.int32, "gate_of_row": torch.float32,
              "xp": torch.bfloat16, "pre": torch.bfloat16, "h": torch.bfloat16, "yo": torch.bfloat16,
              "y_acc": torch.float32, "grid_sync": torch.int32, "ctx_bias": torch.float32}


def _f32c(t: torch.Tensor) -> torch.Tensor:
    """fp32 contiguous view of a parameter (the usual case: returned as it is, detached)."""
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


class _CastJob:
    """The fp32 -> bf16 casts of the expert slabs, deferred so that they can be launched after the router kernel:
    `struct()` for b200moe_layer_fwd (which launches them itself), `fork()` / `launch()` / `join()` for the
    expert-parallel passes that drive the kernels from Python."""

    def __init__(self, mod, a1, a2, side, keep):
        self.mod, self.a1, self.a2, self.side, self.keep = mod, a1, a2, side, keep
        mod.__dict__["_prep_keep"] = keep            # fp32 sources are read on the side stream: alive until the next call
        self._c = None

    def struct(self):
        self._c = _lib.WeightPrep(self.a1[0], self.a1[2], self.a2[0], self.a2[2], self.side.cuda_stream, None)
        return C.addressof(self._c)

    def fork(self):
        self.side.wait_stream(torch.cuda.current_stream(self.side.device))

    def launch(self):
        with torch.cuda.stream(self.side):
            _lib.call("b200moe_prepare_weights", *self.a1, *self.a2, None, None, self.mod.d_model, self.mod.num_experts,
                      _lib.stream_ptr())
            self.ev = torch.cuda.Event()
            self.ev.record(self.side)

    def join(self):
        torch.cuda.current_stream(self.side.device).wait_event(self.ev)


class _LayerPlan:
    """Host-side description of one (T, tokens_per_image) problem: the C descriptor and both workspace layouts."""

    def __init__(self, mod, T: int, tpi: int):
        lib = _lib.load()
        E, k, d, F = mod.num_experts, mod.top_k, mod.d_model, mod.dim_feedforward
        cap = ops.capacity_of(mod.capacity_factor, T, k, E)
        rows_cap = lib.b200moe_layer_rows_cap(T, k, E, cap, d, F)
        if rows_cap <= 0:
            raise RuntimeError(f"b200moe: bad layer shape T={T} k={k} E={E}")
        # mirrors fused_combine() / reduce_combine() of csrc/layer.cu (the workspace layout depends on them)
        self.fused = mod.combine in ("tail", "epilogue") or (mod.combine == "auto" and rows_cap * d * 2 <= (48 << 20))
        self.reduce = self.fused and mod.combine == "epilogue" and k <= 2
        self.gather = mod.dispatch == "gather" and not self.reduce      # gather_dispatch() of csrc/layer.cu
        self.desc = _lib.LayerDesc(T, d, F, E, k, tpi, cap, rows_cap, int(mod.normalize_topk),
                                   mod.aux_loss_weight, mod.z_loss_weight, _lib.ACTIVATIONS[mod.activation],
                                   _lib.COMBINE_FLAGS[mod.combine] | _lib.DISPATCH_FLAGS[mod.dispatch])
        self.desc_ptr = C.addressof(self.desc)
        off = (C.c_int64 * len(_lib.WS_NAMES))()
        self.ws_bytes = lib.b200moe_layer_ws_layout(self.desc_ptr, C.addressof(off))
        boff = (C.c_int64 * len(_lib.BWS_NAMES))()
        self.bws_bytes = lib.b200moe_layer_bwd_ws_layout(self.desc_ptr, C.addressof(boff))
        if self.ws_bytes < 0 or self.bws_bytes < 0:
            raise RuntimeError("b200moe: " + lib.b200moe_last_error().decode())
        self.off = dict(zip(_lib.WS_NAMES, list(off)))
        nblk = lib.b200moe_num_router_blocks(T)
        R = rows_cap
        self.shape = {"logits": (T, E), "topk_idx": (T, k), "topk_gate": (T, k), "block_counts": (nblk, k, E),
                      "block_stats": (nblk, E + 1), "expert_offsets": (E + 1,), "meta": (8,),
                      "tile_expert": (R // 128,), "tokens_per_expert": (E,), "kept_per_expert": (E,), "losses": (3,),
                      "slot_limit": (2 * k * E,), "row_of_assignment": (T, k), "token_of_row": (R,),
                      "gate_of_row": (R,), "xp": (0 if self.gather else R, d), "pre": (R, F), "h": (R, F), "yo": (R, d),
                      "y_acc": (T, d) if self.reduce else (0, d), "grid_sync": (4,) if self.fused else (0,),
                      "ctx_bias": ((T + tpi - 1) // tpi, E)}
        self.capacity, self.rows_cap = cap, rows_cap

    def view(self, ws: torch.Tensor, name: str) -> torch.Tensor:
        dt, shp = _WS_DTYPES[name], self.shape[name]
        n = dt.itemsize * math.prod(shp)
        o = self.off[name]
        return ws[o:o + n].view(dt).view(shp)


class _Aux(Mapping):
    """Telemetry of the last forward (SURVEY.md §8(b) `last_aux`): tensors are views into the forward workspace,
    created on first access so that a training step pays nothing for what it does not read."""
    _ALIAS = {"gates": "topk_gate"}
    _KEYS = ("load_balance", "z_loss", "tokens_per_expert", "kept_per_expert", "dropped", "logits", "topk_idx",
             "gates", "row_of_assignment", "token_of_row", "expert_offsets", "meta", "capacity")

    def __init__(self, plan: _LayerPlan, ws: torch.Tensor):
        self._plan, self._ws, self._cache = plan, ws, {}

    def __getitem__(self, key):
        if key in self._cache:
            return self._cache[key]
        if key not in self._KEYS:
            raise KeyError(key)
        pl, ws = self._plan, self._ws
        if key == "capacity":
            v = pl.capacity
        elif key == "load_balance":
            v = pl.view(ws, "losses")[1]
        elif key == "z_loss":
            v = pl.view(ws, "losses")[2]
        elif key == "dropped":
            v = pl.view(ws, "meta")[2]
        else:
            v = pl.view(ws, self._ALIAS.get(key, key))
        self._cache[key] = v
        return v

    def __iter__(self):
        return iter(self._KEYS)

    def __len__(self):
        return len(self._KEYS)


class _CtxBiasFunction(torch.autograd.Function):
    """A.2 context bias as one kernel each way (used by the expert-parallel path, whose layer pass is not a single
    C call): ctx_bias[B,E] = concat_i Embed_i(context[:, i]) @ Wc^T."""

    @staticmethod
    def forward(ctx, context, wc, mod, *tables):
        B, E = context.shape[0], wc.shape[0]
        tabs = [_f32c(t) for t in tables]
        wcf = _f32c(wc)
        cb = torch.empty((B, E), dtype=torch.float32, device=wc.device)
        tp, vp = mod._ctx_arrays(tabs)
        _lib.call("b200moe_ctx_bias_fwd", context.data_ptr(), tp, vp, len(tabs), mod.context_dim, wcf.data_ptr(), B, E,
                  cb.data_ptr(), _lib.stream_ptr())
        ctx.mod, ctx.saved = mod, (context, tabs, wcf)
        return cb

    @staticmethod
    def backward(ctx, dcb):
        context, tabs, wcf = ctx.saved
        mod = ctx.mod
        dcb = _f32c(dcb)
        dwc = torch.empty_like(wcf)
        dtabs = [torch.empty_like(t) for t in tabs]
        tp, vp = mod._ctx_arrays(tabs)
        dp = (C.c_void_p * len(tabs))(*[t.data_ptr() for t in dtabs])
        _lib.call("b200moe_ctx_bias_bwd", context.data_ptr(), dcb.data_ptr(), tp, vp, len(tabs), mod.context_dim,
                  wcf.data_ptr(), context.shape[0], wcf.shape[0], C.addressof(dp), dwc.data_ptr(), _lib.stream_ptr())
        return (None, dwc, None, *dtabs)


class _MoEFunction(torch.autograd.Function):
    """One C call each way: b200moe_layer_fwd (router -> plan -> permute -> up(+GELU) -> down -> combine) and
    b200moe_layer_bwd (mirror, incl. the A.12 router gradient); the context bias of A.2 and its gradient are one
    more small kernel each."""

    @staticmethod
    def forward(ctx, x2, context, wg, wc, w1, b1, w2, b2, mod, tpi, residual, ln_w, ln_b, ln_eps, *tables):
        T, d = x2.shape
        dev = x2.device
        E = mod.num_experts
        # Function.forward itself always runs with grad mode off, and needs_input_grad is True for parameters even under
        # the caller's torch.no_grad(): the caller's grad mode (noted by ContextMoE.forward) decides what is saved
        need_grad = any(ctx.needs_input_grad) and mod.__dict__.get("_grad_mode", True)
        pl = mod._layer_plan(T, tpi)
        # (while bench.py's per-kernel timing records, the casts stay on the current stream: a cast running next to the
        # router / plan / permute would steal their HBM bandwidth and the numbers would not be those of isolated kernels)
        w1b, w2b, wsplit, wgf, ready = mod._prepared(w1, w2, wg, overlap=not _lib.PROFILING)
        b1f, b2f = _f32c(b1), _f32c(b2)
        st = _lib.stream_ptr()
        cx_ptr = None
        if context is not None and wc is not None:
            # A.2: the bias is evaluated inside the router kernel (per 128-token tile, while its tokens are in flight)
            tabs = [_f32c(t) for t in tables]
            wcf = _f32c(wc)
            cx = mod._ctx_desc(context, tabs, wcf)
            cx_ptr = C.addressof(cx)
            ctx.ctx_saved = (context, tabs, wcf)
        ws = torch.empty((pl.ws_bytes,), dtype=torch.uint8, device=dev)
        pn_ptr = None
        ctx.ln = None
        out_f32 = False
        if residual is not None:                     # fused post-norm: y = LayerNorm(residual + MoE(x))
            # the residual stream keeps its dtype (fp32 under the reference's autocast: LayerNorm outputs fp32) and the
            # normalised output has it too, as with the vendored `norm2(residual + ffn)` (hybrid_encoder.py:158-160)
            out_f32 = residual.dtype == torch.float32
            gam, bet = _f32c(ln_w), _f32c(ln_b)
            mean_rstd = torch.empty((T, 2), dtype=torch.float32, device=dev)
            pn = _lib.PostNorm(residual.data_ptr(), gam.data_ptr(), bet.data_ptr(), float(ln_eps), mean_rstd.data_ptr(),
                               int(out_f32), int(out_f32))
            pn_ptr = C.addressof(pn)
            ctx.ln = (residual, gam, mean_rstd, ln_w.dtype, ln_b.dtype)
        y = torch.empty((T, d), dtype=torch.float32 if out_f32 else torch.bfloat16, device=dev)
        _lib.call("b200moe_layer_fwd", pl.desc_ptr, x2.data_ptr(), None, cx_ptr,
                  wsplit.data_ptr(), w1b.data_ptr(), b1f.data_ptr(), w2b.data_ptr(), b2f.data_ptr(), ws.data_ptr(),
                  y.data_ptr(), int(need_grad), pn_ptr, ready.struct() if ready is not None else None, st)
        ctx.mod, ctx.pl, ctx.has_ctx = mod, pl, cx_ptr is not None
        ctx.dtypes = (wg.dtype, w1.dtype, b1.dtype, w2.dtype, b2.dtype)
        ctx.save_for_backward(x2, wgf, w1b, w2b, ws)
        losses = pl.view(ws, "losses")
        aux = losses[0]
        stats = losses.detach()
        # undefined upstream gradients stay None (the engine would otherwise launch a zero fill per missing one, on the
        # critical path between forward and backward); backward handles dy / d_aux == None itself
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(stats)
        mod.__dict__["_last_ws"] = (pl, ws)
        return y, aux, stats

    @staticmethod
    def backward(ctx, dy, d_aux, _d_stats):
        x2, wgf, w1b, w2b, ws = ctx.saved_tensors
        mod, pl = ctx.mod, ctx.pl
        T, d = x2.shape
        E, F = mod.num_experts, mod.dim_feedforward
        dev = x2.device
        need = ctx.needs_input_grad      # x2, context, wg, wc, w1, b1, w2, b2, mod, tpi, residual, ln_w, ln_b, eps, *tables
        res_f32 = ctx.ln is not None and ctx.ln[0].dtype == torch.float32
        if dy is None:
            dy = torch.zeros_like(x2)
        dy = dy.contiguous().to(torch.float32 if res_f32 else torch.bfloat16)
        d_aux_t = _f32c(d_aux) if d_aux is not None else None
        f32 = torch.float32
        dx = torch.empty((T, d), dtype=torch.bfloat16, device=dev) if need[0] else None
        dwg = torch.empty((E, d), dtype=f32, device=dev) if need[2] else None
        dw1 = torch.empty((E, F, d), dtype=f32, device=dev) if need[4] else None
        db1 = torch.empty((E, F), dtype=f32, device=dev) if need[5] else None
        dw2 = torch.empty((E, d, F), dtype=f32, device=dev) if need[6] else None
        db2 = torch.empty((E, d), dtype=f32, device=dev) if need[7] else None
        want_ctx = ctx.has_ctx and (need[3] or any(need[14:]))
        dwc, dtabs, cx_ptr = None, (), None
        if want_ctx:
            # d(ctx_proj), d(ctx_embed) are written by the same C call, on its router branch
            context, tabs, wcf = ctx.ctx_saved
            dwc = torch.empty_like(wcf)
            flat = torch.empty((sum(t.numel() for t in tabs),), dtype=f32, device=dev)
            dtabs, o = [], 0
            for t in tabs:
                dtabs.append(flat[o:o + t.numel()].view(t.shape))
                o += t.numel()
            cx = mod._ctx_desc(context, tabs, wcf, dtabs, dwc)
            cx_ptr = C.addressof(cx)
            dtabs = tuple(dtabs)
        bws = torch.empty((pl.bws_bytes,), dtype=torch.uint8, device=dev)
        p = lambda t: t.data_ptr() if t is not None else None  # noqa: E731
        st = _lib.stream_ptr()
        pn_ptr, dres, dgam, dbet = None, None, None, None
        if ctx.ln is not None:
            residual, gam, mean_rstd, t_lw, t_lb = ctx.ln
            dv = torch.empty((T, d), dtype=torch.bfloat16, device=dev)        # d(MoE output): feeds the expert backward
            dres = torch.empty((T, d), dtype=f32, device=dev) if res_f32 else dv
            dgam, dbet = torch.empty((d,), dtype=f32, device=dev), torch.empty((d,), dtype=f32, device=dev)
            scr = torch.empty((_lib.load().b200moe_ln_scratch_floats(T, d),), dtype=f32, device=dev)
            pn = _lib.PostNormBwd(residual.data_ptr(), gam.data_ptr(), mean_rstd.data_ptr(), dv.data_ptr(),
                                  scr.data_ptr(), dgam.data_ptr(), dbet.data_ptr(), int(res_f32), int(res_f32),
                                  dres.data_ptr() if res_f32 else None)
            pn_ptr = C.addressof(pn)
        _lib.call("b200moe_layer_bwd", pl.desc_ptr, x2.data_ptr(), wgf.data_ptr(), w1b.data_ptr(), w2b.data_ptr(),
                  ws.data_ptr(), bws.data_ptr(), dy.data_ptr(), p(d_aux_t), p(dx), None, p(dwg), p(dw1), p(db1),
                  p(dw2), p(db2), int(ctx.has_ctx), cx_ptr, pn_ptr, st)
        if ctx.ln is not None:
            dgam, dbet = (dgam if dgam.dtype == t_lw else dgam.to(t_lw)), (dbet if dbet.dtype == t_lb else dbet.to(t_lb))
        t_wg, t_w1, t_b1, t_w2, t_b2 = ctx.dtypes

        def cast(g, dt):
            return g if g is None or g.dtype == dt else g.to(dt)
        n_tab = len(need) - 14
        if not dtabs:
            dtabs = (None,) * n_tab
        return (dx, None, cast(dwg, t_wg), dwc, cast(dw1, t_w1), cast(db1, t_b1), cast(dw2, t_w2), cast(db2, t_b2),
                None, None, dres, dgam, dbet, None, *dtabs)


class ContextMoE(nn.Module):
    """Context-gated top-k MoE FFN; parameters: gate.weight [E,d], ctx_embed.{i}.weight,
    ctx_proj.weight [E, n_fields*context_dim], w1 [E,F,d], b1 [E,F], w2 [E,d,F], b2 [E,d] (under expert parallelism
    the four expert tensors hold this rank's E / world_size experts)."""

    def __init__(self, d_model: int, dim_feedforward: int, num_experts: int, top_k: int = 2,
                 capacity_factor: Optional[float] = 1.25, context_vocab_sizes: Sequence[int] = (),
                 context_dim: int = 32, activation: str = "gelu", aux_loss_weight: float = 1e-2,
                 z_loss_weight: float = 1e-3, normalize_topk: bool = True, ep_group=None,
                 ep_p2p: Optional[bool] = None, ep_max_tokens: Optional[int] = None):
        super().__init__()
        if activation not in _lib.ACTIVATIONS:
            # encoder FFN: exact-erf GELU (zoo/rtdetr/utils.py:160-161); decoder FFN: ReLU (rtdetrv2_decoder.py:169)
            raise ValueError(f"activation must be one of {sorted(_lib.ACTIVATIONS)}, got {activation!r}")
        if not 1 <= top_k <= num_experts:
            raise ValueError("top_k must be in [1, num_experts]")
        if num_experts > 64 or top_k > 8:
            raise ValueError("num_experts <= 64 and top_k <= 8 are supported")
        if d_model % 128 or dim_feedforward % 128:
            raise ValueError("d_model and dim_feedforward must be multiples of 128 (tcgen05 tiles)")
        self.d_model, self.dim_feedforward = d_model, dim_feedforward
        self.num_experts, self.top_k = num_experts, top_k
        self.capacity_factor = capacity_factor
        self.aux_loss_weight, self.z_loss_weight = float(aux_loss_weight), float(z_loss_weight)
        self.normalize_topk = bool(normalize_topk)
        self.context_vocab_sizes = tuple(context_vocab_sizes)
        self.context_dim = context_dim
        self.activation = activation
        # where the A.7 combine runs: "tail": inside the down-projection kernel, after a grid-wide rendezvous, from the expert
        # outputs it just stored; "epilogue": in the down-projection's epilogue itself by fp32 reductions (top_k <= 2;
        # measured slower from ~10^4 tokens on: L2 atomic throughput); "separate": down + combine kernels, the tested
        # reference of both; "auto" (default): tail while yo fits L2, separate beyond (include/b200moe.h has the numbers)
        self.combine = "auto"
        # how the tokens reach the experts: "permute" (default): b200moe_permute writes xp [R,d], the GEMMs read it with
        # tiled TMA loads; "gather": the up-projection's (and, backward, the w1 weight-gradient's) TMA loads fetch
        # x[token_of_row[r]] themselves (tile::gather4), no permuted copy exists -- bit-identical, measured slower on
        # B200 (include/b200moe.h, B200MOE_LAYER_GATHER_DISPATCH)
        self.dispatch = "permute"
        self.ep_group = ep_group
        self.ep_p2p = ep_p2p          # None: peer-memory dispatch when available; False: NCCL all-to-all; True: required
        self.ep_max_tokens = ep_max_tokens   # tokens per rank the peer-memory arena is sized for (None: the first batch)
        # debugging aid (one host sync per forward): raise on context ids outside [0, vocab); by default such ids behave
        # like "missing" without its embedding (they contribute nothing to the bias)
        self.validate_context = False
        self.ep_chunks = 1            # groups of local experts whose all-to-all overlaps the other groups' GEMMs
        # (measured at 2 GPUs, workload D: 1 group 3.72 ms, 2 groups 4.00 ms, 4 groups 4.91 ms per step)
        self.ep_size, self.ep_rank = 1, 0
        if ep_group is not None:
            import torch.distributed as dist
            self.ep_size, self.ep_rank = dist.get_world_size(ep_group), dist.get_rank(ep_group)
        if num_experts % self.ep_size:
            raise ValueError("num_experts must be divisible by the expert-parallel world size")
        self.num_local_experts = num_experts // self.ep_size   # experts [ep_rank*EL, (ep_rank+1)*EL) live here
        # expert-parallel backward multiplies the gradients of w1, b1, w2, b2 by this: 1 / ep_size puts them on DDP's
        # mean-over-ranks scale (b200moe.checkpoint.ddp_ignore_experts sets it; 1.0 = plain sum over all ranks' losses)
        self.expert_grad_scale = 1.0

        E, Fh, d = self.num_local_experts, dim_feedforward, d_model
        self.gate = nn.Linear(d, num_experts, bias=False)
        self.ctx_embed = nn.ModuleList([nn.Embedding(v, context_dim) for v in self.context_vocab_sizes])
        nctx = len(self.context_vocab_sizes) * context_dim
        self.ctx_proj = nn.Linear(nctx, num_experts, bias=False) if nctx else None
        self.w1 = nn.Parameter(torch.empty(E, Fh, d))
        self.b1 = nn.Parameter(torch.empty(E, Fh))
        self.w2 = nn.Parameter(torch.empty(E, d, Fh))
        self.b2 = nn.Parameter(torch.empty(E, d))
        self.reset_parameters()
        self.aux_loss = torch.zeros(())
        self.last_aux = {}
        self._wcache: dict = {}      # bf16 operand copies of the weights, keyed by parameter version
        self._plans: dict = {}       # (T, tokens_per_image) -> _LayerPlan

    # module invariants (§8(b)): deepcopy-safe, plain state_dict -> caches and process groups are not copied
    def __deepcopy__(self, memo):
        import copy
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k_, v in self.__dict__.items():
            if k_ in ("_ep_comm_stream", "_last_ws", "_last", "_ctx_keep", "_ep_static", "_ep_pinned", "_p2p_arena", "_p2p_side", "_p2p_side1", "_prep_stream",
                      "_prep_keep"):
                continue
            if k_ in ("_wcache", "last_aux", "_plans"):
                new.__dict__[k_] = {}
            elif k_ == "aux_loss":
                new.__dict__[k_] = torch.zeros(())
            elif k_ == "ep_group":
                new.__dict__[k_] = v
            else:
                new.__dict__[k_] = copy.deepcopy(v, memo)
        return new

    def __getstate__(self):
        s = dict(self.__dict__)
        s["_wcache"], s["last_aux"], s["aux_loss"], s["ep_group"], s["_plans"] = {}, {}, torch.zeros(()), None, {}
        for k_ in ("_ep_comm_stream", "_last_ws", "_last", "_ctx_keep", "_ep_static", "_ep_pinned", "_p2p_arena", "_p2p_side", "_p2p_side1", "_prep_stream",
                      "_prep_keep"):
            s.pop(k_, None)
        return s

    def reset_parameters(self):
        """A.6: every expert initialised like nn.Linear; gate N(0, d^-1/2); ctx_proj N(0, 0.02)."""
        lo = self.ep_rank * self.num_local_experts
        with torch.no_grad():
            # every rank draws all E experts in the same order and keeps its own slice, so a seed gives the same
            # global model whatever the expert-parallel world size
            for e in range(self.num_experts):
                w1 = nn.init.kaiming_uniform_(torch.empty(self.dim_feedforward, self.d_model), a=math.sqrt(5))
                w2 = nn.init.kaiming_uniform_(torch.empty(self.d_model, self.dim_feedforward), a=math.sqrt(5))
                b1 = torch.empty(self.dim_feedforward).uniform_(-1.0 / math.sqrt(self.d_model),
                                                                1.0 / math.sqrt(self.d_model))
                b2 = torch.empty(self.d_model).uniform_(-1.0 / math.sqrt(self.dim_feedforward),
                                                        1.0 / math.sqrt(self.dim_feedforward))
                if lo <= e < lo + self.num_local_experts:
                    self.w1[e - lo].copy_(w1); self.w2[e - lo].copy_(w2)
                    self.b1[e - lo].copy_(b1); self.b2[e - lo].copy_(b2)
            nn.init.normal_(self.gate.weight, std=self.d_model ** -0.5)
            if self.ctx_proj is not None:
                nn.init.normal_(self.ctx_proj.weight, std=0.02)

    def _stash(self, r, p):
        self._last = (r, p)

    def _layer_plan(self, T: int, tpi: int) -> _LayerPlan:
        # everything the descriptor is built from: a later change of capacity_factor, the loss weights, ... must not
        # be served a stale plan
        key = (T, tpi, self.capacity_factor, self.normalize_topk, self.aux_loss_weight, self.z_loss_weight,
               self.activation, self.combine, self.dispatch)
        pl = self._plans.get(key)
        if pl is None:
            if len(self._plans) > 64:       # multiscale training varies T per step (dataloader.yml:22)
                self._plans.clear()
            pl = self._plans[key] = _LayerPlan(self, T, tpi)
        return pl

    def _prepared(self, w1, w2, wg, overlap: bool = False):
        """bf16 operand copies of w1 / w2 and the bf16 hi/lo split of the gate, redone only when a parameter changed
        (version counters).  Returns (w1_bf16, w2_bf16, wg_split, wg_fp32, job): with `overlap` the two big casts are NOT
        launched here but returned as a `_CastJob` (None when nothing is to do) that the forward runs on a side stream
        next to the router / plan / permute; without it, one launch on the current stream."""
        key = (w1.data_ptr(), w1._version, w2.data_ptr(), w2._version, wg.data_ptr(), wg._version, str(w1.device))
        hit = self._wcache.get("prep")
        if hit is not None and hit[0] == key:
            return (*hit[1], None)
        dev, bf = w1.device, torch.bfloat16
        w1d, w2d = w1.detach().contiguous(), w2.detach().contiguous()
        wgf = wg.detach().float().contiguous()
        w1b = w1d if w1d.dtype == bf else torch.empty(w1d.shape, dtype=bf, device=dev)
        w2b = w2d if w2d.dtype == bf else torch.empty(w2d.shape, dtype=bf, device=dev)
        nsplit = _lib.load().b200moe_router_scratch_bytes(self.d_model, self.num_experts) // 2
        wsplit = torch.empty((nsplit,), dtype=bf, device=dev)
        w1f = w1d.float() if w1d.dtype not in (bf, torch.float32) else w1d
        w2f = w2d.float() if w2d.dtype not in (bf, torch.float32) else w2d
        a1 = (w1f.data_ptr(), w1b.data_ptr(), w1f.numel()) if w1b is not w1d else (None, w1b.data_ptr(), 0)
        a2 = (w2f.data_ptr(), w2b.data_ptr(), w2f.numel()) if w2b is not w2d else (None, w2b.data_ptr(), 0)
        ready = None
        if overlap and (a1[2] or a2[2]):
            side = self.__dict__.get("_prep_stream")
            if side is None or side.device != dev:
                side = self.__dict__["_prep_stream"] = torch.cuda.Stream(device=dev)
            # only the gate split is launched here; the casts of the expert slabs are handed back as a job that the
            # forward launches on the side stream AFTER its router kernel (b200moe_weight_prep in include/b200moe.h has
            # the reason) and joins before the up-projection
            _lib.call("b200moe_prepare_weights", None, w1b.data_ptr(), 0, None, w2b.data_ptr(), 0, wgf.data_ptr(),
                      wsplit.data_ptr(), self.d_model, self.num_experts, _lib.stream_ptr())
            ready = _CastJob(self, a1, a2, side, (w1f, w2f))
        else:
            _lib.call("b200moe_prepare_weights", *a1, *a2, wgf.data_ptr(), wsplit.data_ptr(), self.d_model,
                      self.num_experts, _lib.stream_ptr())
        out = (w1b, w2b, wsplit, wgf)
        self._wcache["prep"] = (key, out)
        return (*out, ready)

    def _ep_call(self, ep_forward, xb, cb, tpi):
        """Expert-parallel pass: over peer memory (b200moe/ep_p2p.py) when the group's GPUs can map each other's memory,
        a capacity factor bounds the rows per rank and no unfinished backward owns the arena; else NCCL all-to-alls."""
        import os
        if (self.ep_p2p is not False and self.capacity_factor is not None and self.ep_chunks == 1
                and os.environ.get("B200MOE_EP_P2P", "1") != "0"):
            from . import ep_p2p
            try:
                ar = ep_p2p.arena_for(self, xb.shape[0], xb.device)
            except ep_p2p.P2PUnavailable:      # raised on every rank or on none (the set-up agrees collectively)
                if self.ep_p2p:
                    raise
                self.ep_p2p, ar = False, None
            if ar is not None and not ar.pending:
                return ep_p2p.p2p_forward(self, xb, cb, tpi, ar)
        return ep_forward(self, xb, cb, tpi)

    def _ctx_desc(self, context, tabs, wcf, dtabs=None, dwc=None):
        """b200moe_ctx_desc (a HOST struct of device pointers) for the whole-layer calls; the caller keeps it alive for
        the duration of the call."""
        n = len(tabs)
        cx = _lib.CtxDesc()
        cx.context = context.data_ptr()
        for i, t in enumerate(tabs):
            cx.tables[i] = t.data_ptr()
            cx.vocab[i] = t.shape[0]
            if dtabs is not None:
                cx.dtables[i] = dtabs[i].data_ptr()
        cx.n_fields, cx.context_dim, cx.B = n, self.context_dim, context.shape[0]
        cx.wc = wcf.data_ptr()
        cx.dwc = dwc.data_ptr() if dwc is not None else None
        return cx

    def _ctx_arrays(self, tabs):
        """HOST arrays (device pointers of the embedding tables, their vocabulary sizes) for the ctx_bias calls."""
        key = tuple(t.data_ptr() for t in tabs)
        keep = self.__dict__.get("_ctx_keep")
        if keep is None or keep[0] != key:
            arr = (C.c_void_p * len(tabs))(*key)
            voc = (C.c_int32 * len(tabs))(*[t.shape[0] for t in tabs])
            keep = self.__dict__["_ctx_keep"] = (key, arr, voc, C.addressof(arr), C.addressof(voc))
        return keep[3], keep[4]

    def forward_post_norm(self, x: torch.Tensor, residual: torch.Tensor, norm: nn.LayerNorm,
                          context: Optional[torch.Tensor] = None) -> torch.Tensor:
        """`norm(residual + self(x, context))` (hybrid_encoder.py:157-160 with dropout 0) with the residual add and the
        LayerNorm fused into the combine kernel, forward and backward.  Returns a tensor of x's shape and dtype."""
        return self.forward(x, context, _post=(residual, norm))

    def forward(self, x: torch.Tensor, context: Optional[torch.Tensor] = None, _post=None) -> torch.Tensor:
        if not x.is_cuda:
            raise RuntimeError("b200moe.ContextMoE runs on a B200 (sm_100a) only; there is no CPU fallback")
        if x.shape[-1] != self.d_model:
            raise ValueError(f"last dim {x.shape[-1]} != d_model {self.d_model}")
        if x.device.index != torch.cuda.current_device():
            # the C calls launch on the CURRENT device's stream: make x's device current for the duration of the call
            with torch.cuda.device(x.device):
                return self.forward(x, context, _post)
        shape = x.shape
        x2 = x.reshape(-1, self.d_model)
        T = x2.shape[0]
        self.__dict__["_grad_mode"] = torch.is_grad_enabled()
        if T == 0:
            self.aux_loss = x.new_zeros((), dtype=torch.float32)
            return x.clone() if _post is None else _post[1](_post[0] + x)
        if _post is not None and (self.ep_size > 1 or self.d_model > 1024 or not isinstance(_post[1], nn.LayerNorm)
                                  or _post[1].weight is None or _post[1].bias is None):
            return _post[1](_post[0] + self.forward(x, context))        # shapes the fused kernel does not cover
        B = shape[0] if x.dim() == 3 else 1
        if context is not None and context.shape[0] != B:
            raise ValueError("context must have one row per image")
        if context is not None and self.validate_context and self.ctx_proj is not None:
            vs = torch.tensor(self.context_vocab_sizes, device=context.device)
            if context.shape[1] != len(self.context_vocab_sizes) or bool(((context < 0) | (context >= vs)).any()):
                raise ValueError(f"context ids outside the vocabulary {self.context_vocab_sizes}: {context.tolist()}")
        tpi = T // B
        xb = x2 if x2.dtype == torch.bfloat16 else x2.to(torch.bfloat16)
        if not xb.is_contiguous():
            xb = xb.contiguous()
        if self.ep_size > 1:
            from .ep import ep_forward
            cb = None
            if context is not None and self.ctx_proj is not None:
                cb = _CtxBiasFunction.apply(context.long().contiguous(), self.ctx_proj.weight, self,
                                            *[m_.weight for m_ in self.ctx_embed])
            y, aux, stats = self._ep_call(ep_forward, xb, cb, tpi)
            self.aux_loss = aux
            r, p = self._last
            self.last_aux = {"load_balance": stats[1], "z_loss": stats[2], "tokens_per_expert": p.tokens_per_expert,
                             "kept_per_expert": p.kept_per_expert, "dropped": p.meta[2], "logits": r.logits,
                             "topk_idx": r.topk_idx, "gates": r.topk_gate, "row_of_assignment": p.row_of_assignment,
                             "token_of_row": p.token_of_row, "expert_offsets": p.expert_offsets, "meta": p.meta,
                             "capacity": p.capacity}
            return y.to(x.dtype).reshape(shape)
        use_ctx = context is not None and self.ctx_proj is not None
        ctx_ids = context.long().contiguous() if use_ctx else None
        tables = [m.weight for m in self.ctx_embed] if use_ctx else []
        res_b = ln_w = ln_b = None
        ln_eps = 0.0
        if _post is not None:
            res, norm = _post
            if res.shape != x.shape:
                raise ValueError("residual must have the shape of x")
            res_b = res.reshape(-1, self.d_model)
            if res_b.dtype not in (torch.bfloat16, torch.float32):
                res_b = res_b.float()
            res_b = res_b if res_b.is_contiguous() else res_b.contiguous()
            ln_w, ln_b, ln_eps = norm.weight, norm.bias, norm.eps
        y, aux, _stats = _MoEFunction.apply(xb, ctx_ids, self.gate.weight,
                                            self.ctx_proj.weight if use_ctx else None, self.w1, self.b1, self.w2,
                                            self.b2, self, tpi, res_b, ln_w, ln_b, ln_eps, *tables)
        # plain attributes: bypass nn.Module.__setattr__ (several microseconds per assignment)
        self.__dict__["aux_loss"] = aux
        self.__dict__["last_aux"] = _Aux(*self._last_ws)
        return (y if y.dtype == x.dtype else y.to(x.dtype)).reshape(shape)
