"""ctypes binding of libb200moe.so (include/b200moe.h).  No CPU fallback: a missing library is a hard error."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200moe.so")

_P, _I, _F, _L = C.c_void_p, C.c_int, C.c_float, C.c_int64

# name -> argtypes (return type is int unless listed in _RESTYPE); mirrors include/b200moe.h one to one
SIGNATURES = {
    "b200moe_version": [],
    "b200moe_launch_count": [_I],
    "b200moe_last_error": [],
    "b200moe_check_device": [],
    "b200moe_rows_upper_bound": [_I, _I, _I, _I],
    "b200moe_layer_row_pad": [_I, _I, _I, _I, _I, _I],
    "b200moe_layer_rows_cap": [_I, _I, _I, _I, _I, _I],
    "b200moe_router_block": [_I],
    "b200moe_num_router_blocks": [_I],
    "b200moe_cast_f32_bf16": [_P, _P, _L, _P],
    "b200moe_router_scratch_bytes": [_I, _I],
    "b200moe_router_fwd": [_P, _P, _I, _P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P],
    "b200moe_router_fwd_ctx": [_P, _P, _I, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P],
    "b200moe_plan": [_P, _P, _P, _P, _I, _I, _I, _I, _P, _P, _I, _I, _F, _F, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "b200moe_permute": [_P, _P, _P, _I, _I, _P, _P],
    "b200moe_permute_init": [_P, _P, _P, _I, _I, _P, _P, _P, _I, _P],
    "b200moe_combine": [_P, _P, _P, _I, _I, _I, _P, _P],
    "b200moe_combine_bwd": [_P, _P, _P, _P, _P, _I, _I, _P, _P, _P],
    "b200moe_ffn_up": [_P, _P, _P, _P, _P, _I, _I, _I, _I, _P, _P, _I, _I, _P],
    "b200moe_ffn_down": [_P, _P, _P, _P, _P, _I, _I, _I, _I, _P, _I, _P],
    "b200moe_ffn_down_combine": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _I, _P],
    "b200moe_ffn_dgrad_down": [_P, _P, _P, _P, _P, _I, _I, _I, _I, _P, _I, _I, _P],
    "b200moe_ffn_dgrad_up": [_P, _P, _P, _P, _I, _I, _I, _I, _P, _I, _P],
    "b200moe_ffn_wgrad": [_P, _P, _P, _I, _I, _I, _I, _P, _I, _P],
    "b200moe_ffn_wgrad2": [_P, _P, _I, _I, _P, _P, _P, _I, _I, _P, _P, _I, _I, _I, _P],
    "b200moe_ffn_up_gather": [_P, _P, _I, _P, _P, _P, _P, _I, _I, _I, _I, _P, _P, _I, _I, _P, _P],
    "b200moe_ffn_wgrad_gather": [_P, _P, _P, _I, _P, _I, _I, _I, _I, _P, _I, _P],
    "b200moe_colsum": [_P, _I, _P, _P, _I, _P, _P, _I, _I, _P],
    "b200moe_prepare_weights": [_P, _P, _L, _P, _P, _L, _P, _P, _I, _I, _P],
    "b200moe_ctx_bias_fwd": [_P, _P, _P, _I, _I, _P, _I, _I, _P, _P],
    "b200moe_ctx_bias_bwd": [_P, _P, _P, _P, _I, _I, _P, _I, _I, _P, _P, _P],
    "b200moe_ep_p2p_tables": [_P, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P],
    "b200moe_ep_p2p_zero_pads": [_P, _P, _I, _I, _I, _I, _P, _P, _P],
    "b200moe_ep_p2p_dispatch": [_P, _P, _P, _P, _P, _P, _L, _I, _I, _I, _I, _I, _P, _P],
    "b200moe_ep_p2p_combine": [_P, _L, _I, _P, _P, _P, _I, _I, _I, _P, _P],
    "b200moe_ep_p2p_combine_bwd": [_P, _P, _L, _L, _I, _P, _P, _P, _P, _I, _I, _P, _P],
    "b200moe_ep_p2p_pull_rows": [_P, _L, _I, _P, _P, _I, _I, _P, _P],
    "b200moe_ep_p2p_unpermute_bwd": [_P, _L, _I, _P, _P, _P, _P, _I, _I, _I, _I, _P, _P],
    "b200moe_layer_profile": [_I],
    "b200moe_layer_pdl": [_I],
    "b200moe_layer_fork": [_I],
    "b200moe_layer_profile_read": [_I, _P, _P],
    "b200moe_layer_ws_layout": [_P, _P],
    "b200moe_layer_bwd_ws_layout": [_P, _P],
    "b200moe_layer_fwd": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _P, _P, _P],
    "b200moe_layer_bwd": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _P, _P, _P],
    "b200moe_combine_ln_fwd": [_P, _P, _P, _P, _P, _I, _P, _P, _F, _I, _I, _I, _P, _I, _P, _P],
    "b200moe_ln_scratch_floats": [_I, _I],
    "b200moe_combine_ln_bwd": [_P, _I, _P, _P, _P, _P, _P, _I, _P, _P, _I, _I, _I, _P, _P, _P, _P, _P, _P],
    "b200moe_router_bwd_dlogits": [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _F, _F, _P, _P],
    "b200moe_dwg_scratch_floats": [_I, _I, _I],
    "b200moe_router_bwd_dwg": [_P, _P, _I, _I, _I, _P, _P, _P],
    "b200moe_router_bwd_dwg_noalloc": [_P, _P, _I, _I, _I, _P, _P],
    "b200moe_router_bwd_dctx": [_P, _I, _I, _I, _P, _P],
    "b200moe_router_bwd_params": [_P, _P, _I, _I, _I, _I, _P, _P, _P, _P],
    "b200moe_unpermute_bwd": [_P, _P, _P, _P, _I, _I, _I, _I, _P, _P],
    "b200moe_ep_recv_plan": [_P, _P, _I, _I, _I, _P, _P, _P, _P, _P],
    "b200moe_scatter_rows": [_P, _P, _P, _I, _I, _P, _P],
    "b200moe_ep_hist": [_P, _I, _I, _I, _I, _P, _P],
    "b200moe_ep_route_tables": [_P, _I, _I, _I, _I, C.c_double, _I, _P, _P, _P, _P],
}
_RESTYPE = {"b200moe_last_error": C.c_char_p, "b200moe_launch_count": C.c_int64, "b200moe_dwg_scratch_floats": C.c_int64,
            "b200moe_router_scratch_bytes": C.c_int64, "b200moe_layer_ws_layout": C.c_int64,
            "b200moe_layer_bwd_ws_layout": C.c_int64, "b200moe_ln_scratch_floats": C.c_int64}


class PostNorm(C.Structure):
    """b200moe_postnorm of include/b200moe.h."""
    _fields_ = [("residual", C.c_void_p), ("gamma", C.c_void_p), ("beta", C.c_void_p), ("eps", C.c_float),
                ("mean_rstd", C.c_void_p), ("residual_f32", C.c_int32), ("out_f32", C.c_int32)]


class PostNormBwd(C.Structure):
    """b200moe_postnorm_bwd of include/b200moe.h."""
    _fields_ = [("residual", C.c_void_p), ("gamma", C.c_void_p), ("mean_rstd", C.c_void_p), ("dv", C.c_void_p),
                ("scratch", C.c_void_p), ("dgamma", C.c_void_p), ("dbeta", C.c_void_p), ("residual_f32", C.c_int32),
                ("dout_f32", C.c_int32), ("dres_f32", C.c_void_p)]


MAX_CTX_FIELDS = 8


class CtxDesc(C.Structure):
    """b200moe_ctx_desc of include/b200moe.h."""
    _fields_ = [("context", C.c_void_p), ("tables", C.c_void_p * MAX_CTX_FIELDS), ("vocab", C.c_int32 * MAX_CTX_FIELDS),
                ("n_fields", C.c_int32), ("context_dim", C.c_int32), ("B", C.c_int32), ("wc", C.c_void_p),
                ("dtables", C.c_void_p * MAX_CTX_FIELDS), ("dwc", C.c_void_p)]


class LayerDesc(C.Structure):
    """b200moe_layer_desc of include/b200moe.h."""
    _fields_ = [("T", C.c_int32), ("d", C.c_int32), ("F", C.c_int32), ("E", C.c_int32), ("k", C.c_int32),
                ("tokens_per_image", C.c_int32), ("capacity", C.c_int32), ("rows_cap", C.c_int32),
                ("normalize_topk", C.c_int32), ("aux_loss_weight", C.c_float), ("z_loss_weight", C.c_float),
                ("activation", C.c_int32), ("flags", C.c_int32)]


class WeightPrep(C.Structure):
    """b200moe_weight_prep (include/b200moe.h)."""
    _fields_ = [("w1_f32", C.c_void_p), ("n1", C.c_int64), ("w2_f32", C.c_void_p), ("n2", C.c_int64),
                ("side_stream", C.c_void_p), ("ready", C.c_void_p)]


ACT_GELU, ACT_RELU = 0, 1                 # B200MOE_ACT_*
ACTIVATIONS = {"gelu": ACT_GELU, "relu": ACT_RELU}
LAYER_SEPARATE_COMBINE = 1                # B200MOE_LAYER_SEPARATE_COMBINE
LAYER_COMBINE_REDUCE = 2                  # B200MOE_LAYER_COMBINE_REDUCE
LAYER_COMBINE_TAIL = 4                    # B200MOE_LAYER_COMBINE_TAIL
LAYER_GATHER_DISPATCH = 8                 # B200MOE_LAYER_GATHER_DISPATCH
DISPATCH_FLAGS = {"permute": 0, "gather": LAYER_GATHER_DISPATCH}
COMBINE_FLAGS = {"auto": 0, "tail": LAYER_COMBINE_TAIL, "epilogue": LAYER_COMBINE_REDUCE,
                 "separate": LAYER_SEPARATE_COMBINE}


WS_NAMES = ("logits", "topk_idx", "topk_gate", "block_counts", "block_stats", "expert_offsets", "meta", "tile_expert",
            "tokens_per_expert", "kept_per_expert", "losses", "slot_limit", "row_of_assignment", "token_of_row",
            "gate_of_row", "xp", "pre", "h", "yo", "y_acc", "grid_sync", "ctx_bias")
BWS_NAMES = ("dyp", "dgate_row", "dhpre", "dxp", "dlogits", "dwg_scratch", "dctx")

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Load the shared library once.  Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python multimodal-moe_b200/build.py` "
            "(b200moe has no CPU or PyTorch fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here == header and library out of sync
        fn.argtypes = argtypes
        fn.restype = _RESTYPE.get(name, C.c_int)
    _lib = lib
    return lib


def ptr(t: Optional[torch.Tensor]):
    """Device pointer of a contiguous tensor (None -> NULL)."""
    if t is None:
        return None
    assert t.is_contiguous(), "b200moe: tensors passed to the C ABI must be contiguous"
    return t.data_ptr()


def stream_ptr() -> int:
    """Raw cudaStream_t of torch's current stream on the current device."""
    return torch._C._cuda_getCurrentRawStream(torch.cuda.current_device())


# kernels launched by each entry point (memset nodes not counted); used for bench.py's gpu_launches
KERNELS_PER_CALL = {
    "b200moe_cast_f32_bf16": 1, "b200moe_prepare_weights": 1, "b200moe_ctx_bias_fwd": 1, "b200moe_ctx_bias_bwd": 1,
    "b200moe_layer_fwd": 6, "b200moe_layer_bwd": 11, "b200moe_router_fwd_ctx": 1, "b200moe_combine_ln_fwd": 1, "b200moe_combine_ln_bwd": 2, "b200moe_router_bwd_params": 2, "b200moe_router_fwd": 2, "b200moe_plan": 2, "b200moe_permute": 1,
    "b200moe_combine": 1, "b200moe_combine_bwd": 1, "b200moe_ffn_up": 1, "b200moe_ffn_down": 1,
    "b200moe_ffn_dgrad_down": 1, "b200moe_ffn_dgrad_up": 1, "b200moe_ffn_wgrad": 1, "b200moe_colsum": 1,
    "b200moe_permute_init": 1, "b200moe_ffn_down_combine": 1, "b200moe_ffn_up_gather": 1, "b200moe_ffn_wgrad2": 1, "b200moe_ffn_wgrad_gather": 1,
    "b200moe_router_bwd_dlogits": 1, "b200moe_router_bwd_dwg": 2, "b200moe_router_bwd_dwg_noalloc": 2, "b200moe_router_bwd_dctx": 1,
    "b200moe_unpermute_bwd": 1, "b200moe_ep_recv_plan": 1, "b200moe_ep_hist": 1, "b200moe_ep_route_tables": 1, "b200moe_scatter_rows": 1,
}


class Counters:
    """Launch counter and optional per-call CUDA-event timing (bench.py's instrumented pass)."""
    launches = 0
    timing = False
    events: list = []   # (name, start_event, end_event)

    @classmethod
    def reset(cls, timing: bool = False):
        cls.launches, cls.timing, cls.events = 0, timing, []

    @classmethod
    def durations_ms(cls) -> dict:
        """name -> list of per-call durations (call after torch.cuda.synchronize())."""
        out: dict = {}
        for name, a, b in cls.events:
            out.setdefault(name, []).append(a.elapsed_time(b))
        return out


PROFILING = False      # per-kernel event timing is recording: the layer keeps every kernel on one stream meanwhile


def layer_profile(enable: bool) -> None:
    global PROFILING
    PROFILING = bool(enable)
    load().b200moe_layer_profile(int(enable))


def layer_profile_durations() -> dict:
    """step name -> list of durations (ms) recorded inside b200moe_layer_fwd / _bwd since layer_profile(True);
    call after torch.cuda.synchronize()."""
    lib, out, i = load(), {}, 0
    name, ms = C.c_char_p(), C.c_float()
    while lib.b200moe_layer_profile_read(i, C.byref(name), C.byref(ms)) == 0:
        out.setdefault(name.value.decode(), []).append(ms.value)
        i += 1
    return out


def call(name: str, *args) -> None:
    lib = load()
    if Counters.timing:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rc = getattr(lib, name)(*args)
        b.record()
        Counters.events.append((name, a, b))
    else:
        rc = getattr(lib, name)(*args)
    Counters.launches += KERNELS_PER_CALL.get(name, 0)
    if rc != 0:
        msg = lib.b200moe_last_error()
        raise RuntimeError(f"{name} failed ({rc}): {msg.decode() if msg else ''}")
