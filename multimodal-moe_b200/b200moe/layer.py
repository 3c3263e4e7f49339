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

import math
from typing import Optional, Sequence

import torch
import torch.nn as nn

from . import ops


class _MoEFunction(torch.autograd.Function):
    """fwd: router -> plan -> permute -> up(+GELU) -> down -> combine;  bwd mirrors it (A.12 router gradient)."""

    @staticmethod
    def forward(ctx, x2, ctx_bias, wg, w1, b1, w2, b2, mod, tokens_per_image):
        T, d = x2.shape
        E, k = mod.num_experts, mod.top_k
        need_grad = any(ctx.needs_input_grad)   # forward runs under no_grad: ask autograd, not the grad mode
        w1b, w2b = mod._bf16_weight("w1", w1), mod._bf16_weight("w2", w2)
        b1f, b2f = b1.float().contiguous(), b2.float().contiguous()
        wgf = wg.float().contiguous()
        cb = ctx_bias.float().contiguous() if ctx_bias is not None else None

        r = ops.router_fwd(x2, wgf, cb, tokens_per_image, k, mod.normalize_topk)
        cap = ops.capacity_of(mod.capacity_factor, T, k, E)
        p = ops.plan(r, T, k, E, cap, mod.aux_loss_weight, mod.z_loss_weight)
        xp = ops.permute(x2, p.token_of_row, p.meta, p.rows_cap)
        pre, h = ops.ffn_up(xp, w1b, b1f, p.meta, p.tile_expert, save_pre=need_grad)
        yo = ops.ffn_down(h, w2b, b2f, p.meta, p.tile_expert)
        y = ops.combine(yo, p.row_of_assignment, r.topk_gate)

        ctx.mod, ctx.tpi, ctx.plan, ctx.routing = mod, tokens_per_image, p, r
        ctx.has_ctx = ctx_bias is not None
        ctx.dtypes = (w1.dtype, b1.dtype, w2.dtype, b2.dtype, wg.dtype)
        ctx.save_for_backward(x2, wgf, w1b, w2b, xp, pre, h, yo)
        aux = p.losses[0]
        stats = p.losses.detach()
        ctx.mark_non_differentiable(stats)
        mod._stash(r, p)
        return y, aux, stats

    @staticmethod
    def backward(ctx, dy, d_aux, _d_stats):
        x2, wgf, w1b, w2b, xp, pre, h, yo = ctx.saved_tensors
        mod, p, r = ctx.mod, ctx.plan, ctx.routing
        T, d = x2.shape
        E, k = mod.num_experts, mod.top_k
        need = ctx.needs_input_grad
        if dy is None:
            dy = torch.zeros_like(x2)
        dy = dy.contiguous().to(torch.bfloat16)
        d_aux_t = d_aux.float().contiguous() if d_aux is not None else None

        dyp, dgate_row = ops.combine_bwd(dy, yo, p.token_of_row, p.gate_of_row, p.meta, p.rows_cap)
        dhpre = ops.ffn_dgrad_down(dyp, w2b, pre, p.meta, p.tile_expert)
        dw2 = db2 = dw1 = db1 = None
        if need[5]:
            dw2 = ops.ffn_wgrad(dyp, h, p.expert_offsets, E)
        if need[6]:
            db2 = ops.colsum(dyp, p.expert_offsets, E)
        if need[3]:
            dw1 = ops.ffn_wgrad(dhpre, xp, p.expert_offsets, E)
        if need[4]:
            db1 = ops.colsum(dhpre, p.expert_offsets, E)
        dlogits = ops.router_bwd_dlogits(r.logits, r.topk_idx, p.row_of_assignment, dgate_row, p.tokens_per_expert,
                                         d_aux_t, mod.normalize_topk, mod.aux_loss_weight, mod.z_loss_weight)
        dx = None
        if need[0]:
            dxp = ops.ffn_dgrad_up(dhpre, w1b, p.meta, p.tile_expert)
            dx = ops.unpermute_bwd(dxp, p.row_of_assignment, dlogits, wgf, T, k, d)
        dwg = ops.router_bwd_dwg(dlogits, x2) if need[2] else None
        dctx = ops.router_bwd_dctx(dlogits, ctx.tpi) if (ctx.has_ctx and need[1]) else None
        t_w1, t_b1, t_w2, t_b2, t_wg = ctx.dtypes

        def cast(g, dt):
            return g if g is None or g.dtype == dt else g.to(dt)
        return (dx, dctx, cast(dwg, t_wg), cast(dw1, t_w1), cast(db1, t_b1), cast(dw2, t_w2), cast(db2, t_b2),
                None, None)


class ContextMoE(nn.Module):
    """Context-gated top-k MoE FFN; parameters: gate.weight [E,d], ctx_embed.{i}.weight,
    ctx_proj.weight [E, n_fields*context_dim], w1 [E,F,d], b1 [E,F], w2 [E,d,F], b2 [E,d] (under expert parallelism
    the four expert tensors hold this rank's E / world_size experts)."""

    def __init__(self, d_model: int, dim_feedforward: int, num_experts: int, top_k: int = 2,
                 capacity_factor: Optional[float] = 1.25, context_vocab_sizes: Sequence[int] = (),
                 context_dim: int = 32, activation: str = "gelu", aux_loss_weight: float = 1e-2,
                 z_loss_weight: float = 1e-3, normalize_topk: bool = True, ep_group=None):
        super().__init__()
        if activation != "gelu":
            raise ValueError("only exact-erf GELU is defined for this layer (zoo/rtdetr/utils.py:160-161)")
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
        self.ep_group = ep_group
        self.ep_chunks = 2            # groups of local experts whose all-to-all overlaps the other groups' GEMMs
        self.ep_size, self.ep_rank = 1, 0
        if ep_group is not None:
            import torch.distributed as dist
            self.ep_size, self.ep_rank = dist.get_world_size(ep_group), dist.get_rank(ep_group)
        if num_experts % self.ep_size:
            raise ValueError("num_experts must be divisible by the expert-parallel world size")
        self.num_local_experts = num_experts // self.ep_size   # experts [ep_rank*EL, (ep_rank+1)*EL) live here

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
        self.last_aux: dict = {}
        self._wcache: dict = {}

    # module invariants (§8(b)): deepcopy-safe, plain state_dict -> caches and process groups are not copied
    def __deepcopy__(self, memo):
        import copy
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k_, v in self.__dict__.items():
            if k_ == "_ep_comm_stream":
                continue
            if k_ in ("_wcache", "last_aux"):
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
        s["_wcache"], s["last_aux"], s["aux_loss"], s["ep_group"] = {}, {}, torch.zeros(()), None
        s.pop("_ep_comm_stream", None)
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

    def _bf16_weight(self, name: str, w: torch.Tensor) -> torch.Tensor:
        """bf16 operand copy of a master weight, re-cast only when the parameter changed (version counter)."""
        if w.dtype == torch.bfloat16:
            return w.detach().contiguous()
        key = (w.data_ptr(), w._version, tuple(w.shape))
        hit = self._wcache.get(name)
        if hit is not None and hit[0] == key:
            return hit[1]
        wb = ops.cast_bf16(w.detach())
        self._wcache[name] = (key, wb)
        return wb

    def _stash(self, r, p):
        self._last = (r, p)

    def context_bias(self, context: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        """A.2: ctx_bias[B,E] = concat_i Embed_i(context[:, i]) @ Wc^T (fp32 regardless of autocast)."""
        if context is None or self.ctx_proj is None:
            return None
        with torch.autocast(device_type=context.device.type, enabled=False):
            emb = torch.cat([m(context[:, i]) for i, m in enumerate(self.ctx_embed)], dim=-1)
            return torch.nn.functional.linear(emb.float(), self.ctx_proj.weight.float())

    def forward(self, x: torch.Tensor, context: Optional[torch.Tensor] = None) -> torch.Tensor:
        if not x.is_cuda:
            raise RuntimeError("b200moe.ContextMoE runs on a B200 (sm_100a) only; there is no CPU fallback")
        if x.shape[-1] != self.d_model:
            raise ValueError(f"last dim {x.shape[-1]} != d_model {self.d_model}")
        shape = x.shape
        x2 = x.reshape(-1, self.d_model)
        T = x2.shape[0]
        if T == 0:
            self.aux_loss = x.new_zeros((), dtype=torch.float32)
            return x.clone()
        B = shape[0] if x.dim() == 3 else 1
        if context is not None and context.shape[0] != B:
            raise ValueError("context must have one row per image")
        tpi = T // B
        xb = x2.to(torch.bfloat16).contiguous()
        cb = self.context_bias(context)
        if self.ep_size > 1:
            from .ep import ep_forward
            y, aux, stats = ep_forward(self, xb, cb, tpi)
        else:
            y, aux, stats = _MoEFunction.apply(xb, cb, self.gate.weight, self.w1, self.b1, self.w2, self.b2, self, tpi)
        self.aux_loss = aux
        r, p = self._last
        self.last_aux = {"load_balance": stats[1], "z_loss": stats[2], "tokens_per_expert": p.tokens_per_expert,
                         "kept_per_expert": p.kept_per_expert, "dropped": p.meta[2], "logits": r.logits,
                         "topk_idx": r.topk_idx, "gates": r.topk_gate, "row_of_assignment": p.row_of_assignment,
                         "token_of_row": p.token_of_row, "expert_offsets": p.expert_offsets, "meta": p.meta,
                         "capacity": p.capacity}
        return y.to(x.dtype).reshape(shape)
