"""Tensor-level wrappers over the C ABI (include/b200moe.h).  torch is used only for device memory and streams."""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr

ROW_PAD = 128
ROUTER_BLOCK = 128


def capacity_of(capacity_factor: Optional[float], T: int, k: int, E: int) -> int:
    """capacity = ceil(cf * T * k / E); None => unlimited (T*k)."""
    if capacity_factor is None:
        return T * k
    return int(math.ceil(capacity_factor * T * k / E))


def rows_upper_bound(T: int, k: int, E: int, capacity: int) -> int:
    r = _lib.load().b200moe_rows_upper_bound(T, k, E, capacity)
    if r < 0:
        raise RuntimeError(f"rows_upper_bound({T},{k},{E},{capacity}) failed")
    return r


def _empty(shape, dtype, dev):
    return torch.empty(shape, dtype=dtype, device=dev)


def cast_bf16(src: torch.Tensor) -> torch.Tensor:
    """fp32 -> bf16 operand copy (bf16 inputs are returned as they are)."""
    if src.dtype == torch.bfloat16:
        return src.contiguous()
    src = src.contiguous().float()
    dst = _empty(src.shape, torch.bfloat16, src.device)
    call("b200moe_cast_f32_bf16", ptr(src), ptr(dst), src.numel(), stream_ptr())
    return dst


@dataclass
class Routing:
    logits: torch.Tensor        # [T,E] fp32
    topk_idx: torch.Tensor      # [T,k] int32
    topk_gate: torch.Tensor     # [T,k] fp32
    block_counts: torch.Tensor  # [nblk,k,E] int32 (overwritten by plan with row bases)
    block_stats: torch.Tensor   # [nblk,E+1] fp32


def router_fwd(x: torch.Tensor, wg: Optional[torch.Tensor], ctx_bias: Optional[torch.Tensor],
               tokens_per_image: int, k: int, normalize: bool, wg_split: Optional[torch.Tensor] = None,
               E: Optional[int] = None) -> Routing:
    """wg [E,d] fp32, split into bf16 hi/lo inside the call; or wg=None with wg_split from prepare_weights (and E)."""
    T, d = x.shape
    E = wg.shape[0] if wg is not None else E
    dev = x.device
    nblk = _lib.load().b200moe_num_router_blocks(T)
    r = Routing(_empty((T, E), torch.float32, dev), _empty((T, k), torch.int32, dev),
                _empty((T, k), torch.float32, dev), _empty((nblk, k, E), torch.int32, dev),
                _empty((nblk, E + 1), torch.float32, dev))
    nscratch = _lib.load().b200moe_router_scratch_bytes(d, E)
    if nscratch < 0:
        raise RuntimeError(f"router_fwd: num_experts={E} outside [1,64]")
    wsplit = wg_split if wg_split is not None else _empty((nscratch // 2,), torch.bfloat16, dev)
    call("b200moe_router_fwd", ptr(x), ptr(ctx_bias), tokens_per_image, ptr(wg), ptr(wsplit), T, d, E, k,
         int(normalize),
         ptr(r.logits), ptr(r.topk_idx), ptr(r.topk_gate), ptr(r.block_counts), ptr(r.block_stats), stream_ptr())
    return r


@dataclass
class Plan:
    rows_cap: int
    capacity: int
    expert_offsets: torch.Tensor     # [E+1] int32
    meta: torch.Tensor               # [8] int32
    tile_expert: torch.Tensor        # [rows_cap/128] int32
    tokens_per_expert: torch.Tensor  # [E] int32 (pre-drop)
    kept_per_expert: torch.Tensor    # [E] int32
    losses: torch.Tensor             # [3] fp32: aux, lb, z
    row_of_assignment: torch.Tensor  # [T,k] int32
    token_of_row: torch.Tensor       # [rows_cap] int32
    gate_of_row: torch.Tensor        # [rows_cap] fp32


def plan(r: Routing, T: int, k: int, E: int, capacity: int, aux_w: float, z_w: float,
         global_offset: Optional[torch.Tensor] = None, expert_order: Optional[torch.Tensor] = None,
         row_pad: int = ROW_PAD, rows_cap: Optional[int] = None) -> Plan:
    dev = r.topk_idx.device
    if rows_cap is None:
        rows_cap = rows_upper_bound(T, k, E, capacity)
    i32 = torch.int32
    p = Plan(rows_cap, capacity, _empty((E + 1,), i32, dev), _empty((8,), i32, dev),
             _empty((rows_cap // ROW_PAD,), i32, dev), _empty((E,), i32, dev), _empty((E,), i32, dev),
             _empty((3,), torch.float32, dev), _empty((T, k), i32, dev), _empty((rows_cap,), i32, dev),
             _empty((rows_cap,), torch.float32, dev))
    slot_limit = _empty((2 * k * E,), i32, dev)
    call("b200moe_plan", ptr(r.topk_idx), ptr(r.topk_gate), ptr(r.block_counts), ptr(r.block_stats), T, k, E,
         capacity, ptr(global_offset), ptr(expert_order), row_pad, rows_cap, aux_w, z_w, ptr(p.expert_offsets),
         ptr(p.meta), ptr(p.tile_expert), ptr(p.tokens_per_expert), ptr(p.kept_per_expert), ptr(p.losses),
         ptr(slot_limit), ptr(p.row_of_assignment), ptr(p.token_of_row), ptr(p.gate_of_row), stream_ptr())
    return p


def permute(x: torch.Tensor, token_of_row: torch.Tensor, meta: torch.Tensor, rows_cap: int) -> torch.Tensor:
    d = x.shape[1]
    xp = _empty((rows_cap, d), torch.bfloat16, x.device)
    call("b200moe_permute", ptr(x), ptr(token_of_row), ptr(meta), rows_cap, d, ptr(xp), stream_ptr())
    return xp


def scatter_rows(src: torch.Tensor, src_of_row: torch.Tensor, meta: torch.Tensor, out_rows: int) -> torch.Tensor:
    rows_cap, d = src.shape
    out = _empty((out_rows, d), torch.bfloat16, src.device)
    call("b200moe_scatter_rows", ptr(src), ptr(src_of_row), ptr(meta), rows_cap, d, ptr(out), stream_ptr())
    return out


def combine(yo: torch.Tensor, row_of_assignment: torch.Tensor, topk_gate: torch.Tensor) -> torch.Tensor:
    T, k = row_of_assignment.shape
    d = yo.shape[1]
    y = _empty((T, d), torch.bfloat16, yo.device)
    call("b200moe_combine", ptr(yo), ptr(row_of_assignment), ptr(topk_gate), T, k, d, ptr(y), stream_ptr())
    return y


def combine_bwd(dy: torch.Tensor, yo: Optional[torch.Tensor], token_of_row: torch.Tensor, gate_of_row: torch.Tensor,
                meta: torch.Tensor, rows_cap: int):
    d = dy.shape[1]
    dyp = _empty((rows_cap, d), torch.bfloat16, dy.device)
    dgate_row = _empty((rows_cap,), torch.float32, dy.device) if yo is not None else None
    call("b200moe_combine_bwd", ptr(dy), ptr(yo), ptr(token_of_row), ptr(gate_of_row), ptr(meta), rows_cap, d,
         ptr(dyp), ptr(dgate_row), stream_ptr())
    return dyp, dgate_row


def ffn_up(xp, w1, b1, meta, tile_expert, save_pre: bool = True, row_pad: int = ROW_PAD,
           activation: int = _lib.ACT_GELU):
    rows_cap, d = xp.shape
    E, F, _ = w1.shape
    pre = _empty((rows_cap, F), torch.bfloat16, xp.device) if save_pre else None
    h = _empty((rows_cap, F), torch.bfloat16, xp.device)
    call("b200moe_ffn_up", ptr(xp), ptr(w1), ptr(b1), ptr(meta), ptr(tile_expert), rows_cap, d, F, E, ptr(pre),
         ptr(h), activation, row_pad, stream_ptr())
    return pre, h


def ffn_up_gather(x, token_of_row, w1, b1, meta, tile_expert, rows_cap: int, save_pre: bool = True,
                  row_pad: int = ROW_PAD, activation: int = _lib.ACT_GELU, clear: Optional[torch.Tensor] = None):
    """ffn_up on rows x[token_of_row[r]] gathered by the kernel's TMA loads (no permuted copy)."""
    T, d = x.shape
    E, F, _ = w1.shape
    pre = _empty((rows_cap, F), torch.bfloat16, x.device) if save_pre else None
    h = _empty((rows_cap, F), torch.bfloat16, x.device)
    call("b200moe_ffn_up_gather", ptr(x), ptr(token_of_row), T, ptr(w1), ptr(b1), ptr(meta), ptr(tile_expert), rows_cap,
         d, F, E, ptr(pre), ptr(h), activation, row_pad, ptr(clear), stream_ptr())
    return pre, h


def permute_init(x: torch.Tensor, token_of_row: torch.Tensor, meta: torch.Tensor, rows_cap: int, finalize: bool = True,
                 reduce: bool = True):
    """Dispatch + the zero state of the fused combine: (xp, y_acc [T,d] fp32 | None, grid_sync [4] int32 | None)."""
    T, d = x.shape
    xp = _empty((rows_cap, d), torch.bfloat16, x.device)
    y_acc = _empty((T, d), torch.float32, x.device) if reduce else None
    sync = _empty((4,), torch.int32, x.device) if finalize else None
    call("b200moe_permute_init", ptr(x), ptr(token_of_row), ptr(meta), rows_cap, d, ptr(xp), ptr(y_acc), ptr(sync), T,
         stream_ptr())
    return xp, y_acc, sync


def ffn_down_combine(h, w2, b2, meta, tile_expert, token_of_row, gate_of_row, row_of_assignment, topk_gate,
                     y_acc, grid_sync, row_pad: int = ROW_PAD, save_yo: bool = True):
    """Down-projection + combine in one kernel.  y_acc None: gather mode (the tail combines from yo); else reduction mode
    (top_k <= 2, the epilogue adds into y_acc).  Returns (yo | None, y | None)."""
    rows_cap, F = h.shape
    E, d, _ = w2.shape
    T, k = row_of_assignment.shape
    yo = _empty((rows_cap, d), torch.bfloat16, h.device) if (save_yo or y_acc is None) else None
    y = _empty((T, d), torch.bfloat16, h.device) if grid_sync is not None else None
    call("b200moe_ffn_down_combine", ptr(h), ptr(w2), ptr(b2), ptr(meta), ptr(tile_expert), ptr(token_of_row),
         ptr(gate_of_row), ptr(row_of_assignment), ptr(topk_gate), T, k, rows_cap, d, F, E, ptr(yo), ptr(y_acc),
         ptr(grid_sync), ptr(y), None, row_pad, stream_ptr())
    return yo, y


def ffn_down(h, w2, b2, meta, tile_expert, row_pad: int = ROW_PAD, out: Optional[torch.Tensor] = None):
    rows_cap, F = h.shape
    E, d, _ = w2.shape
    yo = out if out is not None else _empty((rows_cap, d), torch.bfloat16, h.device)
    call("b200moe_ffn_down", ptr(h), ptr(w2), ptr(b2), ptr(meta), ptr(tile_expert), rows_cap, d, F, E, ptr(yo),
         row_pad, stream_ptr())
    return yo


def ffn_dgrad_down(dyp, w2, pre, meta, tile_expert, row_pad: int = ROW_PAD, activation: int = _lib.ACT_GELU):
    rows_cap, d = dyp.shape
    E, _, F = w2.shape
    dhpre = _empty((rows_cap, F), torch.bfloat16, dyp.device)
    call("b200moe_ffn_dgrad_down", ptr(dyp), ptr(w2), ptr(pre), ptr(meta), ptr(tile_expert), rows_cap, d, F, E,
         ptr(dhpre), activation, row_pad, stream_ptr())
    return dhpre


def ffn_dgrad_up(dhpre, w1, meta, tile_expert, row_pad: int = ROW_PAD, out: Optional[torch.Tensor] = None):
    rows_cap, F = dhpre.shape
    E, _, d = w1.shape
    dxp = out if out is not None else _empty((rows_cap, d), torch.bfloat16, dhpre.device)
    call("b200moe_ffn_dgrad_up", ptr(dhpre), ptr(w1), ptr(meta), ptr(tile_expert), rows_cap, d, F, E, ptr(dxp),
         row_pad, stream_ptr())
    return dxp


def ffn_wgrad(a, b, expert_offsets, E: int, out: Optional[torch.Tensor] = None, accumulate: bool = False):
    rows_cap, M = a.shape
    N = b.shape[1]
    if out is None:
        out = _empty((E, M, N), torch.float32, a.device)
        accumulate = False
    call("b200moe_ffn_wgrad", ptr(a), ptr(b), ptr(expert_offsets), rows_cap, M, N, E, ptr(out), int(accumulate),
         stream_ptr())
    return out


def ffn_wgrad2(a0, b0, a1, b1, expert_offsets, E: int, out0: Optional[torch.Tensor] = None,
               out1: Optional[torch.Tensor] = None, accumulate: bool = False):
    """Both weight gradients in one launch: (a0^T b0 [E,M0,N0], a1^T b1 [E,M1,N1]) per expert."""
    rows_cap, M0 = a0.shape
    N0, M1, N1 = b0.shape[1], a1.shape[1], b1.shape[1]
    if out0 is None or out1 is None:
        out0 = _empty((E, M0, N0), torch.float32, a0.device)
        out1 = _empty((E, M1, N1), torch.float32, a0.device)
        accumulate = False
    call("b200moe_ffn_wgrad2", ptr(a0), ptr(b0), M0, N0, ptr(out0), ptr(a1), ptr(b1), M1, N1, ptr(out1),
         ptr(expert_offsets), rows_cap, E, int(accumulate), stream_ptr())
    return out0, out1


def ffn_wgrad_gather(a, x, token_of_row, expert_offsets, E: int, out: Optional[torch.Tensor] = None,
                     accumulate: bool = False):
    """ffn_wgrad with b = x[token_of_row[r]] gathered by the kernel's TMA loads."""
    rows_cap, M = a.shape
    T, N = x.shape
    if out is None:
        out = _empty((E, M, N), torch.float32, a.device)
        accumulate = False
    call("b200moe_ffn_wgrad_gather", ptr(a), ptr(x), ptr(token_of_row), T, ptr(expert_offsets), rows_cap, M, N, E,
         ptr(out), int(accumulate), stream_ptr())
    return out


def colsum(a, expert_offsets, E: int, out: Optional[torch.Tensor] = None, accumulate: bool = False,
           a1: Optional[torch.Tensor] = None, out1: Optional[torch.Tensor] = None):
    """db[e,:] = column sums of a over the rows of expert e (and the same for a1 in the same launch)."""
    n = a.shape[1]
    if out is None:
        out = _empty((E, n), torch.float32, a.device)
        accumulate = False
    n1 = 0
    if a1 is not None:
        n1 = a1.shape[1]
        if out1 is None:
            out1 = _empty((E, n1), torch.float32, a.device)
    call("b200moe_colsum", ptr(a), n, ptr(out), ptr(a1), n1, ptr(out1), ptr(expert_offsets), E, int(accumulate),
         stream_ptr())
    return out if a1 is None else (out, out1)


def router_bwd_dlogits(logits, topk_idx, row_of_assignment, dgate_row, tokens_per_expert, d_aux, normalize: bool,
                       aux_w: float, z_w: float):
    T, E = logits.shape
    k = topk_idx.shape[1]
    dl = _empty((T, E), torch.float32, logits.device)
    call("b200moe_router_bwd_dlogits", ptr(logits), ptr(topk_idx), ptr(row_of_assignment), ptr(dgate_row),
         ptr(tokens_per_expert), ptr(d_aux), T, E, k, int(normalize), aux_w, z_w, ptr(dl), stream_ptr())
    return dl


def router_bwd_dwg(dlogits, x):
    T, E = dlogits.shape
    d = x.shape[1]
    n = _lib.load().b200moe_dwg_scratch_floats(T, d, E)
    scratch = _empty((n,), torch.float32, x.device)
    dwg = _empty((E, d), torch.float32, x.device)
    call("b200moe_router_bwd_dwg", ptr(dlogits), ptr(x), T, d, E, ptr(scratch), ptr(dwg), stream_ptr())
    return dwg


def router_bwd_params(dlogits, x, tokens_per_image: int, want_dwg: bool = True, want_dctx: bool = True):
    """(dwg [E,d] or None, dctx_bias [B,E] or None) in two launches."""
    T, E = dlogits.shape
    d = x.shape[1]
    dev = x.device
    dwg = scratch = dctx = None
    if want_dwg:
        n = _lib.load().b200moe_dwg_scratch_floats(T, d, E)
        scratch = _empty((n,), torch.float32, dev)
        dwg = _empty((E, d), torch.float32, dev)
    if want_dctx:
        B = (T + tokens_per_image - 1) // tokens_per_image
        dctx = _empty((B, E), torch.float32, dev)
    if want_dwg or want_dctx:
        call("b200moe_router_bwd_params", ptr(dlogits), ptr(x), T, d, E, tokens_per_image, ptr(scratch), ptr(dwg),
             ptr(dctx), stream_ptr())
    return dwg, dctx


def router_bwd_dctx(dlogits, tokens_per_image: int):
    T, E = dlogits.shape
    B = (T + tokens_per_image - 1) // tokens_per_image
    out = _empty((B, E), torch.float32, dlogits.device)
    call("b200moe_router_bwd_dctx", ptr(dlogits), T, E, tokens_per_image, ptr(out), stream_ptr())
    return out


def unpermute_bwd(dxp, row_of_assignment, dlogits, wg, T: int, k: int, d: int):
    E = wg.shape[0]
    dev = wg.device
    dx = _empty((T, d), torch.bfloat16, dev)
    call("b200moe_unpermute_bwd", ptr(dxp), ptr(row_of_assignment), ptr(dlogits), ptr(wg), T, k, d, E, ptr(dx),
         stream_ptr())
    return dx


def ep_recv_plan(recv_counts: torch.Tensor, seg_expert: torch.Tensor, E_local: int, rows_cap: int):
    dev = recv_counts.device
    i32 = torch.int32
    offs, meta = _empty((E_local + 1,), i32, dev), _empty((8,), i32, dev)
    tile_expert, src = _empty((rows_cap // ROW_PAD,), i32, dev), _empty((rows_cap,), i32, dev)
    call("b200moe_ep_recv_plan", ptr(recv_counts), ptr(seg_expert), recv_counts.numel(), E_local, rows_cap,
         ptr(offs), ptr(meta), ptr(tile_expert), ptr(src), stream_ptr())
    return offs, meta, tile_expert, src
