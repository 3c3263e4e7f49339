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
  * EMA needs nothing: it is elementwise over state_dict entries, so each rank averages its own shard.
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
    names = expert_param_names(model)
    prev = list(getattr(model, "_ddp_params_and_buffers_to_ignore", []))
    model._ddp_params_and_buffers_to_ignore = prev + [n for n in names if n not in prev]
    return names


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
