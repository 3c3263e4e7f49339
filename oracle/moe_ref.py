"""CPU/fp32 oracle for the context-gated MoE FFN layer.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this file.  The product path
(``multimodal-moe_b200/b200moe``) never does and has no CPU fallback.

PARITY UNPINNED: the reference repository (scaleoutsystems/multimodal-MoE) contains
no MoE layer, router or expert code, so there is no reference
implementation, golden vector or known-answer test to pin this oracle against.
What *is* pinned (tests/test_oracle.py, tests/golden/):
  * identical experts + renormalised top-k  ==  the vendored dense encoder FFN
    ``linear2(gelu(linear1(x)))`` of
    third_party/rtdetr/rtdetrv2_pytorch/src/zoo/rtdetr/hybrid_encoder.py:129-131,157
    (golden vectors generated from that file by tests/golden/make_golden.py);
  * exact-erf GELU as selected by .../zoo/rtdetr/utils.py:160-161;
  * context vocabulary sizes from scripts/add_solar_context_bins.py:90-107 and
    outputs/analysis/camera/detection/context_field_frequencies_final.csv:2-26;
  * closed forms: uniform router => load_balance == 1, z_loss == ln(E)^2.

Semantics follow SURVEY.md Appendix A (A.1 .. A.12); each function cites its rule.
Everything is plain PyTorch fp32 (stable sort, cumsum, index_select, per-expert
F.linear loop) so autograd provides the reference gradients (A.12).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

# Context vocabulary (A.2, id 0 == "missing").  Sizes INCLUDE the missing id.
# solar: scripts/add_solar_context_bins.py:91-97 (5 labels) + "missing" (:107)
# others: context_field_frequencies_final.csv:2-26 category lists, sorted by name.
CONTEXT_FIELDS = ("solar_context_bin", "scraped_weather", "time_of_day", "road_type", "road_condition")
CONTEXT_VOCAB = {
    "solar_context_bin": ["missing", "night(<-6)", "twilight(-6..0)", "low_sun(0..15)", "mid_sun(15..45)", "high_sun(>45)"],
    "scraped_weather": ["missing"] + sorted(["partly-cloudy-day", "cloudy", "clear-day", "rain", "partly-cloudy-night",
                                             "clear-night", "fog", "snow", "wind"]),
    "time_of_day": ["missing"] + sorted(["day", "night", "twilight"]),
    "road_type": ["missing"] + sorted(["city", "arterial-urban", "highway", "arterial-rural", "smaller-rural"]),
    "road_condition": ["missing"] + sorted(["normal", "wet", "snow"]),
}
CONTEXT_VOCAB_SIZES = tuple(len(CONTEXT_VOCAB[f]) for f in CONTEXT_FIELDS)  # (6, 10, 4, 6, 4)


def capacity_of(capacity_factor: Optional[float], T: int, k: int, E: int) -> int:
    """A.5: capacity = ceil(cf * T * k / E); None => unlimited (returned as T*k)."""
    if capacity_factor is None:
        return T * k
    return int(math.ceil(capacity_factor * T * k / E))


def route(logits: torch.Tensor, top_k: int, normalize_topk: bool = True):
    """A.3/A.4: fp32 softmax, top-k by (p desc, expert index asc), optional renormalisation.

    The ranking is taken on the logits: softmax is monotone, so (logit desc, index asc) IS the (p desc, index asc)
    order of A.4, minus the spurious ties that fp32 rounding of p can fabricate between two distinct logits
    (which would make "bit-exact routing on identical fp32 logits" depend on the exp implementation).
    Returns (p [T,E], idx [T,k] int64, gates [T,k]).
    """
    logits = logits.float()
    p = torch.softmax(logits, dim=-1)
    # stable descending sort == lowest index wins ties; torch.topk does not guarantee that
    idx = torch.sort(logits, dim=-1, descending=True, stable=True).indices[:, :top_k]
    g = torch.gather(p, 1, idx)
    if normalize_topk:
        g = g / (g.sum(dim=-1, keepdim=True) + 1e-9)
    return p, idx, g


def assign_slots(idx: torch.Tensor, num_experts: int, capacity: int,
                 global_offset: Optional[torch.Tensor] = None):
    """A.5: slot-major, then token-index priority; keep iff position-in-expert < capacity.

    idx [T,k] int64.  global_offset [k,E] = number of earlier-priority assignments owned by
    other ranks (expert parallelism); None on a single rank.
    Returns pos [T,k] (position in expert, global order) and keep [T,k] bool.
    """
    T, k = idx.shape
    flat = idx.t().reshape(-1)                                   # slot-major order (s, t)
    onehot = F.one_hot(flat, num_experts).to(torch.int64)        # [k*T, E]
    if global_offset is None:
        before = onehot.cumsum(0) - onehot                       # exclusive count over the whole list
        pos = before.gather(1, flat[:, None]).squeeze(1)
    else:
        oh = onehot.view(k, T, num_experts)
        before = oh.cumsum(1) - oh                               # exclusive within each slot level
        before = before + global_offset.to(torch.int64)[:, None, :]
        pos = before.reshape(k * T, num_experts).gather(1, flat[:, None]).squeeze(1)
    pos = pos.view(k, T).t().contiguous()
    keep = pos < capacity
    return pos, keep


def permutation(idx: torch.Tensor, keep: torch.Tensor, num_experts: int, pad_to: int = 1):
    """Row layout of the permuted buffer (what the CUDA plan kernel must reproduce bit-exactly).

    Kept assignments of expert e occupy rows [off[e], off[e]+kept[e]) in priority order
    (slot-major, then token index); every segment starts at a multiple of ``pad_to``.
    Returns (offsets [E+1], kept [E], slot_of_assignment [T,k] (-1 = dropped), token_of_row [rows] (-1 = pad)).
    """
    T, k = idx.shape
    flat_e = idx.t().reshape(-1)
    flat_keep = keep.t().reshape(-1)
    kept = torch.zeros(num_experts, dtype=torch.int64)
    kept.scatter_add_(0, flat_e[flat_keep], torch.ones(int(flat_keep.sum()), dtype=torch.int64))
    padded = (kept + pad_to - 1) // pad_to * pad_to
    offsets = torch.zeros(num_experts + 1, dtype=torch.int64)
    offsets[1:] = padded.cumsum(0)
    onehot = F.one_hot(flat_e, num_experts).to(torch.int64) * flat_keep[:, None].to(torch.int64)
    rank = (onehot.cumsum(0) - onehot).gather(1, flat_e[:, None]).squeeze(1)
    row = offsets[flat_e] + rank
    row = torch.where(flat_keep, row, torch.full_like(row, -1))
    slot_of_assignment = row.view(k, T).t().contiguous()
    token_of_row = torch.full((int(offsets[-1]),), -1, dtype=torch.int64)
    tok = torch.arange(T).repeat(k)
    token_of_row[row[flat_keep]] = tok[flat_keep]
    return offsets, kept, slot_of_assignment, token_of_row


def aux_losses(logits: torch.Tensor, p: torch.Tensor, idx: torch.Tensor, num_experts: int):
    """A.8 load-balance (Switch/ST-MoE form, f pre-drop, non-differentiable) and A.9 router z-loss."""
    T, k = idx.shape
    counts = torch.bincount(idx.reshape(-1), minlength=num_experts)
    f = counts.to(torch.float32) / float(T * k)
    P = p.mean(dim=0)
    lb = num_experts * torch.sum(f.detach() * P)
    z = torch.logsumexp(logits.float(), dim=-1).pow(2).mean()
    return lb, z, counts


class ContextMoERef(nn.Module):
    """fp32 oracle of the layer (SURVEY.md §8(b) constructor/forward, Appendix A semantics)."""

    def __init__(self, d_model: int, dim_feedforward: int, num_experts: int, top_k: int = 2,
                 capacity_factor: Optional[float] = 1.25, context_vocab_sizes: Sequence[int] = (),
                 context_dim: int = 32, activation: str = "gelu", aux_loss_weight: float = 1e-2,
                 z_loss_weight: float = 1e-3, normalize_topk: bool = True, ep_group=None):
        super().__init__()
        if activation not in ("gelu", "relu"):
            # 'gelu' -> nn.GELU() exact erf (zoo/rtdetr/utils.py:160-161), the encoder FFN of hybrid_encoder.py:157;
            # 'relu' -> nn.ReLU (utils.py:152-153), the decoder FFN of rtdetrv2_decoder.py:169,187,203
            raise ValueError("activation must be 'gelu' or 'relu'")
        self.activation = activation
        if top_k > num_experts:
            raise ValueError("top_k > num_experts")
        self.d_model, self.dim_feedforward = d_model, dim_feedforward
        self.num_experts, self.top_k = num_experts, top_k
        self.capacity_factor = capacity_factor
        self.aux_loss_weight, self.z_loss_weight = aux_loss_weight, z_loss_weight
        self.normalize_topk = normalize_topk
        self.context_vocab_sizes = tuple(context_vocab_sizes)
        self.context_dim = context_dim

        E, Fh, d = num_experts, dim_feedforward, d_model
        self.gate = nn.Linear(d, E, bias=False)                      # A.2: shallow linear gate, no bias
        self.ctx_embed = nn.ModuleList([nn.Embedding(v, context_dim) for v in self.context_vocab_sizes])
        nctx = len(self.context_vocab_sizes) * context_dim
        self.ctx_proj = nn.Linear(nctx, E, bias=False) if nctx else None
        self.w1 = nn.Parameter(torch.empty(E, Fh, d))
        self.b1 = nn.Parameter(torch.empty(E, Fh))
        self.w2 = nn.Parameter(torch.empty(E, d, Fh))
        self.b2 = nn.Parameter(torch.empty(E, d))
        self.reset_parameters()
        self.aux_loss = torch.zeros(())
        self.last_aux: dict = {}

    def reset_parameters(self):
        """A.6: each expert initialised like nn.Linear (kaiming_uniform a=sqrt(5), bias U(+-1/sqrt(fan_in)))."""
        E = self.num_experts
        with torch.no_grad():
            for e in range(E):
                nn.init.kaiming_uniform_(self.w1[e], a=math.sqrt(5))
                nn.init.kaiming_uniform_(self.w2[e], a=math.sqrt(5))
                b = 1.0 / math.sqrt(self.d_model)
                self.b1[e].uniform_(-b, b)
                b = 1.0 / math.sqrt(self.dim_feedforward)
                self.b2[e].uniform_(-b, b)
            nn.init.normal_(self.gate.weight, std=self.d_model ** -0.5)
            if self.ctx_proj is not None:
                nn.init.normal_(self.ctx_proj.weight, std=0.02)

    # -- pieces, exposed so tests can feed the CUDA kernel's own logits through the oracle's routing --
    def context_bias(self, context: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        """A.2: ctx_bias[B,E] = concat_i Embed_i(context[:, i]) @ Wc^T ; None context => no bias."""
        if context is None or self.ctx_proj is None:
            return None
        emb = torch.cat([m(context[:, i]) for i, m in enumerate(self.ctx_embed)], dim=-1)
        return self.ctx_proj(emb.float())

    def logits(self, x2: torch.Tensor, ctx_bias: Optional[torch.Tensor], tokens_per_image: int) -> torch.Tensor:
        lg = F.linear(x2.float(), self.gate.weight.float())
        if ctx_bias is not None:
            lg = lg + ctx_bias.repeat_interleave(tokens_per_image, dim=0)   # A.1: tok2img[t] = t // HW
        return lg

    def experts(self, xe: torch.Tensor, e: int) -> torch.Tensor:
        """A.6: W2_e . gelu_erf(W1_e . x + b1_e) + b2_e  (hybrid_encoder.py:157, dropout 0)."""
        pre = F.linear(xe, self.w1[e].float(), self.b1[e].float())
        h = F.gelu(pre) if self.activation == "gelu" else F.relu(pre)
        return F.linear(h, self.w2[e].float(), self.b2[e].float())

    def forward(self, x: torch.Tensor, context: Optional[torch.Tensor] = None,
                logits_override: Optional[torch.Tensor] = None, logits_value: Optional[torch.Tensor] = None,
                global_offset: Optional[torch.Tensor] = None, capacity_override: Optional[int] = None):
        shape = x.shape
        d, E, k = self.d_model, self.num_experts, self.top_k
        x2 = x.reshape(-1, d).float()
        T = x2.shape[0]
        B = shape[0] if x.dim() == 3 else 1
        hw = T // B
        lg = self.logits(x2, self.context_bias(context), hw) if logits_override is None else logits_override.float()
        if logits_value is not None:
            # parity "on identical fp32 logits" (A.11): take the VALUES from the kernel, keep this graph for autograd
            lg = lg + (logits_value.float() - lg).detach()
        p, idx, g = route(lg, k, self.normalize_topk)
        cap = capacity_override if capacity_override is not None else capacity_of(self.capacity_factor, T, k, E)
        pos, keep = assign_slots(idx, E, cap, global_offset)
        lb, z, counts = aux_losses(lg, p, idx, E)

        # A.7: y[t] = sum over kept slots, in slot order, fp32
        y = torch.zeros_like(x2)
        for s in range(k):
            ys = torch.zeros_like(x2)
            for e in range(E):
                sel = torch.nonzero((idx[:, s] == e) & keep[:, s]).squeeze(1)
                if sel.numel():
                    ys = ys.index_add(0, sel, self.experts(x2.index_select(0, sel), e) * g[sel, s:s + 1])
            y = y + ys
        self.aux_loss = self.aux_loss_weight * lb + self.z_loss_weight * z          # A.10
        self.last_aux = {"load_balance": lb.detach(), "z_loss": z.detach(), "tokens_per_expert": counts,
                         "dropped": int((~keep).sum()), "logits": lg.detach(), "topk_idx": idx, "gates": g.detach(),
                         "keep": keep, "pos": pos, "capacity": cap}
        return y.to(x.dtype).reshape(shape)


def dense_ffn(x: torch.Tensor, w1: torch.Tensor, b1: torch.Tensor, w2: torch.Tensor, b2: torch.Tensor) -> torch.Tensor:
    """Restatement of the vendored dense encoder FFN: linear2(dropout0(gelu(linear1(x))))
    (hybrid_encoder.py:129-131,157; activation = nn.GELU() exact erf, utils.py:160-161)."""
    return F.linear(F.gelu(F.linear(x, w1, b1)), w2, b2)
