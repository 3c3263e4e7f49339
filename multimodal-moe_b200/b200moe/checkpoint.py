"""Expert-parallel checkpoint / DDP / EMA glue.

Under expert parallelism `w1, b1, w2, b2` of a ContextMoE hold this rank's E / world_size experts, everything else is
replicated.  The reference's solver saves `model.state_dict()` on rank 0 only
(third_party/rtdetr/rtdetrv2_pytorch/src/solver/det_solver.py:58-64 via src/misc/dist_utils.py `save_on_master`), wraps
the model in DDP (dist_utils.py:125-147) and keeps an EMA copy that iterates the state dict (src/optim/ema.py:44-53).
  * full_state_dict(model): every rank calls it; expert tensors are all-gathered into their global [E, ...] shape, so
    what rank 0 saves loads into any world size (including 1).
  * load_full_state_dict(model, state): the inverse — every rank slices its own experts out of the global tensors.
  * ddp_ignore_experts(model): names DDP must not all-reduce (call before DDP(model), dist_utils.py:140).  Replicated
    parameters (gate, ctx_embed, ctx_proj) stay in DDP's buckets.
  * THE GRADIENT-SCALE RULE.  DDP hands every replicated parameter the MEAN over the ranks of d(loss_r)/d(theta).  An
    expert shard receives rows from every rank's tokens, so its local gradient is already the SUM over the ranks of
    d(loss_r)/d(theta_e) -- W times the scale of everything else (SURVEY.md §7 hard part 6).  ddp_ignore_experts therefore
    also sets `moe.expert_grad_scale = 1 / ep_size`; the expert-parallel backward multiplies the four expert gradients
    by it, so one learning rate / weight decay / clip threshold means the same thing for every parameter and an
    EP run follows the same trajectory as DDP over the same global batch.
  * clip_grad_norm_(model, max_norm): `torch.nn.utils.clip_grad_norm_(model.parameters(), 0.1)` (det_engine.py:56) sees a
    different expert shard on every rank, hence a different total norm, a different clip coefficient and silently
    diverging replicated weights.  This one all-reduces the expert shards' squared norms over ep_group so that every rank
    computes the norm of the GLOBAL parameter vector and applies one coefficient.
  * EPGradScaler: GradScaler whose inf/nan verdict (unscale_, det_engine.py:55) is all-reduced over ep_group for the same
    reason: one rank's shard overflowing must skip the step on every rank.
  * EMA (src/optim/ema.py:31,44-53) needs nothing: it is elementwise over state_dict entries, so each rank averages its
    own shard; save it with full_state_dict(ema.module) exactly like the model.
  * Where to call full_state_dict: `DetSolver.fit` builds the checkpoint with `self.state_dict()` on every rank and only
    the write is rank-0 (`dist_utils.save_on_master`, det_solver.py:58-64), but `state_dict()` of an expert-parallel model is
    the LOCAL shard: replace `self.state_dict()` there by a dict whose 'model' / 'ema' entries come from full_state_dict
    (a collective: every rank must reach it, so it belongs BEFORE the is-main-process check).
All collectives go through `torch.distributed` on the module's `ep_group` (NCCL on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Dict, List

import torch
import torch.distributed as dist
import torch.nn as nn

from .layer import ContextMoE

EXPERT_PARAMS = ("w1", "b1", "w2", "b2")


def _ep_modules(model: nn.Module):
    return [(name, m) for name, m in model.named_modules() if isinstance(m, ContextMoE)]


def expert_param_names(model: nn.Module) -> List[str]:
    """Qualified names of the rank-local expert tensors of every expert-parallel ContextMoE in `model`."""
    out = []
    for name, m in _ep_modules(model):
        if m.ep_size > 1:
            out += [f"{name}.{p}" if name else p for p in EXPERT_PARAMS]
    return out


def ddp_ignore_experts(model: nn.Module) -> List[str]:
    """Call before `DDP(model)`: the sharded expert tensors leave DDP's buckets and their gradients are scaled to DDP's
    mean-over-ranks convention (see the module docstring)."""
    names = expert_param_names(model)
    prev = list(getattr(model, "_ddp_params_and_buffers_to_ignore", []))
    # DDP matches f"{module_name}.{param_name}": for parameters of the ROOT module that is ".w1", not "w1"
    spelled = names + ["." + n for n in names if "." not in n]
    model._ddp_params_and_buffers_to_ignore = prev + [n for n in spelled if n not in prev]
    for _, m in _ep_modules(model):
        if m.ep_size > 1:
            m.expert_grad_scale = 1.0 / m.ep_size
    return names


def _unwrap(model: nn.Module) -> nn.Module:
    return model.module if hasattr(model, "module") and isinstance(model.module, nn.Module) else model


def clip_grad_norm_(model: nn.Module, max_norm: float, norm_type: float = 2.0) -> torch.Tensor:
    """Drop-in for `torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm)` (det_engine.py:56) that is correct
    under expert parallelism: the norm is that of the global gradient (replicated parameters once, every rank's expert
    shard once), identical on all ranks, so the clipped replicated weights stay bit-identical across ranks.
    Collective over the ep_group(s) of the model's expert-parallel layers; on a model without any it is the torch call."""
    if norm_type != 2.0:
        raise ValueError("clip_grad_norm_: only the 2-norm is implemented (what the reference uses)")
    model = _unwrap(model)
    sharded = {}
    for _, m in _ep_modules(model):
        if m.ep_size > 1:
            for p in EXPERT_PARAMS:
                sharded[id(getattr(m, p))] = m.ep_group
    grads = [p for p in model.parameters() if p.grad is not None]
    if not grads:
        return torch.zeros(())
    dev = grads[0].grad.device
    local = torch.zeros((), dtype=torch.float32, device=dev)
    by_group: Dict[int, list] = {}
    for p in grads:
        sq = p.grad.detach().float().pow(2).sum()
        if id(p) in sharded:
            g = sharded[id(p)]
            by_group.setdefault(id(g), [g, torch.zeros((), dtype=torch.float32, device=dev)])[1].add_(sq)
        else:
            local.add_(sq)
    total = local
    for g, sq in by_group.values():
        dist.all_reduce(sq, group=g)              # every shard of the expert tensors, once each
        total = total + sq
    norm = total.sqrt()
    coef = torch.clamp(max_norm / (norm + 1e-6), max=1.0)
    for p in grads:
        p.grad.detach().mul_(coef.to(p.grad.device))
    return norm


class EPGradScaler(torch.amp.GradScaler):
    """GradScaler for models with expert-parallel layers (the vendored loop: scaler.scale(loss).backward();
    scaler.unscale_(optimizer); clip; scaler.step(optimizer); det_engine.py:52-59).  After the per-rank inf/nan scan of
    unscale_ the verdict is all-reduced (MAX) over `ep_group`, as torch's ShardedGradScaler does for FSDP: an overflow
    in ONE rank's expert shard skips the optimizer step on EVERY rank, so replicated weights and the loss scale do not
    drift apart."""

    def __init__(self, ep_group=None, device: str = "cuda", **kwargs):
        super().__init__(device, **kwargs)
        self._ep_group = ep_group

    def _unscale_grads_(self, optimizer, inv_scale, found_inf, allow_fp16):
        out = super()._unscale_grads_(optimizer, inv_scale, found_inf, allow_fp16)
        if dist.is_available() and dist.is_initialized():
            for t in out.values():
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self._ep_group)
        return out


def full_state_dict(model: nn.Module) -> Dict[str, torch.Tensor]:
    """state_dict with every sharded expert tensor gathered to its global shape (collective: call on all ranks)."""
    state = dict(model.state_dict())
    for name, m in _ep_modules(model):
        if m.ep_size == 1:
            continue
        for p in EXPERT_PARAMS:
            key = f"{name}.{p}" if name else p
            local = state[key].contiguous()
            parts = [torch.empty_like(local) for _ in range(m.ep_size)]
            dist.all_gather(parts, local, group=m.ep_group)
            state[key] = torch.cat(parts, dim=0)          # experts are sharded contiguously: rank r owns [r*EL, (r+1)*EL)
    return state


def load_full_state_dict(model: nn.Module, state: Dict[str, torch.Tensor], strict: bool = True):
    """Load a global-shape state dict (from full_state_dict or from a single-GPU run) into a possibly sharded model."""
    state = dict(state)
    for name, m in _ep_modules(model):
        if m.ep_size == 1:
            continue
        lo, hi = m.ep_rank * m.num_local_experts, (m.ep_rank + 1) * m.num_local_experts
        for p in EXPERT_PARAMS:
            key = f"{name}.{p}" if name else p
            if key in state and state[key].shape[0] == m.num_experts:
                state[key] = state[key][lo:hi]
    return model.load_state_dict(state, strict=strict)
