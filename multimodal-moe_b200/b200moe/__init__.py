"""b200moe — B200-native context-gated MoE FFN layer (host side of libb200moe.so).

Public API: ContextMoE (drop-in nn.Module), MoEEncoderLayer / upcycle_encoder (encoder-layer glue), context (scene-context
vocabulary and wire-through), checkpoint (expert-parallel state dicts), ops (tensor-level wrappers of the C ABI),
load_library().
"""
from ._lib import LIB_PATH, SIGNATURES, load as load_library  # noqa: F401
from . import ops  # noqa: F401
from .layer import ContextMoE  # noqa: F401
from .encoder import MoEEncoderLayer, upcycle_encoder  # noqa: F401
from .graphed import GraphedMoE, GraphedStep, MultiShapeGraphedMoE, PipelinedGraphedStep  # noqa: F401
from . import checkpoint, context, registry  # noqa: F401
from .context import (attach_context, batch_context, collect_aux_loss, context_ids, install_context_hook,  # noqa: F401
                      router_telemetry)
from .checkpoint import (EPGradScaler, clip_grad_norm_, ddp_ignore_experts, expert_param_names,  # noqa: F401
                         full_state_dict, load_full_state_dict)

__all__ = ["ContextMoE", "GraphedMoE", "GraphedStep", "MultiShapeGraphedMoE", "PipelinedGraphedStep", "MoEEncoderLayer", "upcycle_encoder", "context", "checkpoint", "registry", "ops", "load_library",
           "LIB_PATH", "SIGNATURES", "attach_context", "batch_context", "collect_aux_loss", "context_ids",
           "install_context_hook", "router_telemetry", "ddp_ignore_experts", "clip_grad_norm_", "EPGradScaler", "expert_param_names", "full_state_dict",
           "load_full_state_dict"]
