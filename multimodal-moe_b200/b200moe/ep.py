"""Expert parallelism: experts sharded contiguously over the ranks of one NVSwitch box, tokens
stay data-parallel, one token all-to-all each way per pass (NCCL over NVLink).

Routing is independent of the world size by construction (A.5): every rank all-gathers the [k,E] assignment
histograms, derives the same global capacity and its own slot-major / rank-major offsets, so the drops are exactly
those of a single-GPU run on the concatenated batch.

Data flow per forward (rank r, E_local = E / W experts [r*E_local, (r+1)*E_local)):
  router (local) -> hist all_gather -> plan(global_offset, row_pad=1) = send layout, experts back to back
  -> permute into the send buffer -> all_to_all_single (uneven splits) -> ep_recv_plan + permute = 128-row-padded
  expert-major layout -> grouped GEMMs on the local experts -> scatter back to arrival order -> all_to_all_single
  -> combine on the token's home rank.  Backward mirrors it.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from . import ops


def route_tables(hist_all: torch.Tensor, tokens_all: torch.Tensor, rank: int, capacity_factor: Optional[float],
                 top_k: int) -> Tuple[torch.Tensor, int, torch.Tensor]:
    """Pure integer host/device logic shared by every rank (tested on CPU with gloo).

    hist_all [W,k,E]: pre-drop assignments per (rank, slot, expert); tokens_all [W]: tokens per rank.
    Returns (global_offset [k,E] for `rank`, capacity, kept [W,E] = rows every rank sends to every expert).
    """
    W, k, E = hist_all.shape
    hist_all = hist_all.to(torch.int64)
    T_global = int(tokens_all.sum())
    capacity = ops.capacity_of(capacity_factor, T_global, top_k, E)
    tot = hist_all.sum(0)                                   # [k,E]
    lower_slots = tot.cumsum(0) - tot                       # all ranks, earlier slot levels
    before_rank = hist_all.cumsum(0) - hist_all             # same slot level, lower ranks  [W,k,E]
    base = lower_slots[None] + before_rank                  # earlier-priority assignments   [W,k,E]
    room = (capacity - base).clamp_(min=0)
    kept = torch.minimum(hist_all, room).sum(1)             # [W,E]
    return base[rank].to(torch.int32).contiguous(), capacity, kept


def split_sizes(kept: torch.Tensor, rank: int, W: int) -> Tuple[List[int], List[int], torch.Tensor, torch.Tensor]:
    """kept [W,E] (host).  Returns (send_splits[W], recv_splits[W], recv_counts [W*E_local], seg_expert)."""
    E = kept.shape[1]
    EL = E // W
    send = kept[rank].view(W, EL).sum(1).tolist()           # rows this rank sends to each destination rank
    mine = kept[:, rank * EL:(rank + 1) * EL]               # [W, EL] rows arriving from each source, per local expert
    recv = mine.sum(1).tolist()
    recv_counts = mine.reshape(-1).to(torch.int32)          # arrival order: source rank major, local expert minor
    seg_expert = torch.arange(EL, dtype=torch.int32).repeat(W)
    return send, recv, recv_counts, seg_expert


def _a2a(buf: torch.Tensor, out_rows: int, in_splits: List[int], out_splits: List[int], group) -> torch.Tensor:
    out = torch.empty((out_rows, buf.shape[1]), dtype=buf.dtype, device=buf.device)
    dist.all_to_all_single(out, buf, output_split_sizes=out_splits, input_split_sizes=in_splits, group=group)
    return out


class _EPFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x2, ctx_bias, wg, w1, b1, w2, b2, mod, tokens_per_image):
        group = mod.ep_group
        W, rank = dist.get_world_size(group), dist.get_rank(group)
        T, d = x2.shape
        E, k, EL = mod.num_experts, mod.top_k, mod.num_local_experts
        need_grad = any(ctx.needs_input_grad)
        w1b, w2b = mod._bf16_weight("w1", w1), mod._bf16_weight("w2", w2)
        b1f, b2f = b1.float().contiguous(), b2.float().contiguous()
        wgf = wg.float().contiguous()
        cb = ctx_bias.float().contiguous() if ctx_bias is not None else None
        dev = x2.device

        r = ops.router_fwd(x2, wgf, cb, tokens_per_image, k, mod.normalize_topk)
        # histograms of every rank (+ token counts) in one small all-gather
        mine = torch.cat([r.block_counts.sum(0, dtype=torch.int32).reshape(-1),
                          torch.tensor([T], dtype=torch.int32, device=dev)])
        allh = torch.empty((W, mine.numel()), dtype=torch.int32, device=dev)
        dist.all_gather_into_tensor(allh, mine, group=group)
        hist_all, tokens_all = allh[:, :-1].reshape(W, k, E), allh[:, -1]
        goff, cap, kept = route_tables(hist_all, tokens_all, rank, mod.capacity_factor, k)
        kept_h = kept.cpu()                                  # the one host sync of the pass: all-to-all split sizes
        send_splits, recv_splits, recv_counts, seg_expert = split_sizes(kept_h, rank, W)
        n_send, n_recv = sum(send_splits), sum(recv_splits)

        send_cap = ops.rows_upper_bound(T, k, E, -1)
        p = ops.plan(r, T, k, E, cap, mod.aux_loss_weight, mod.z_loss_weight, global_offset=goff.to(dev),
                     row_pad=1, rows_cap=send_cap)
        xs = ops.permute(x2, p.token_of_row, p.meta, send_cap)
        xr = _a2a(xs[:n_send], n_recv, send_splits, recv_splits, group)

        recv_cap = max(128, (n_recv + EL * 127 + 127) // 128 * 128)
        offs, meta_r, tile_e, src_of_row = ops.ep_recv_plan(recv_counts.to(dev), seg_expert.to(dev), EL, recv_cap)
        if n_recv == 0:
            xr = torch.zeros((1, d), dtype=torch.bfloat16, device=dev)
        xp = ops.permute(xr, src_of_row, meta_r, recv_cap)
        pre, h = ops.ffn_up(xp, w1b, b1f, meta_r, tile_e, save_pre=need_grad)
        yo = ops.ffn_down(h, w2b, b2f, meta_r, tile_e)
        yr = ops.scatter_rows(yo, src_of_row, meta_r, max(n_recv, 1))
        ys = _a2a(yr[:n_recv], n_send, recv_splits, send_splits, group)
        yo_home = ys                                         # rows come back in this rank's send layout
        y = ops.combine(yo_home, p.row_of_assignment, r.topk_gate)

        ctx.mod, ctx.tpi, ctx.plan, ctx.routing = mod, tokens_per_image, p, r
        ctx.has_ctx = ctx_bias is not None
        ctx.dtypes = (w1.dtype, b1.dtype, w2.dtype, b2.dtype, wg.dtype)
        ctx.ep = (send_splits, recv_splits, n_send, n_recv, send_cap, recv_cap, offs, meta_r, tile_e, src_of_row)
        ctx.save_for_backward(x2, wgf, w1b, w2b, xp, pre, h, yo_home)
        aux = p.losses[0]
        stats = p.losses.detach()
        ctx.mark_non_differentiable(stats)
        mod._stash(r, p)
        return y, aux, stats

    @staticmethod
    def backward(ctx, dy, d_aux, _d_stats):
        x2, wgf, w1b, w2b, xp, pre, h, yo_home = ctx.saved_tensors
        mod, p, r = ctx.mod, ctx.plan, ctx.routing
        group = mod.ep_group
        send_splits, recv_splits, n_send, n_recv, send_cap, recv_cap, offs, meta_r, tile_e, src_of_row = ctx.ep
        T, d = x2.shape
        k, EL = mod.top_k, mod.num_local_experts
        need = ctx.needs_input_grad
        dev = x2.device
        if dy is None:
            dy = torch.zeros_like(x2)
        dy = dy.contiguous().to(torch.bfloat16)
        d_aux_t = d_aux.float().contiguous() if d_aux is not None else None

        dys, dgate_row = ops.combine_bwd(dy, yo_home, p.token_of_row, p.gate_of_row, p.meta, send_cap)
        dyr = _a2a(dys[:n_send], n_recv, send_splits, recv_splits, group)
        if n_recv == 0:
            dyr = torch.zeros((1, d), dtype=torch.bfloat16, device=dev)
        dyp = ops.permute(dyr, src_of_row, meta_r, recv_cap)
        dhpre = ops.ffn_dgrad_down(dyp, w2b, pre, meta_r, tile_e)
        dw2 = ops.ffn_wgrad(dyp, h, offs, EL) if need[5] else None
        db2 = ops.colsum(dyp, offs, EL) if need[6] else None
        dw1 = ops.ffn_wgrad(dhpre, xp, offs, EL) if need[3] else None
        db1 = ops.colsum(dhpre, offs, EL) if need[4] else None
        dlogits = ops.router_bwd_dlogits(r.logits, r.topk_idx, p.row_of_assignment, dgate_row, p.tokens_per_expert,
                                         d_aux_t, mod.normalize_topk, mod.aux_loss_weight, mod.z_loss_weight)
        dx = None
        if need[0]:
            dxp = ops.ffn_dgrad_up(dhpre, w1b, meta_r, tile_e)
            dxr = ops.scatter_rows(dxp, src_of_row, meta_r, max(n_recv, 1))
            dxs = _a2a(dxr[:n_recv], n_send, recv_splits, send_splits, group)
            dx = ops.unpermute_bwd(dxs, p.row_of_assignment, dlogits, wgf, T, k, d)
        dwg = ops.router_bwd_dwg(dlogits, x2) if need[2] else None
        dctx = ops.router_bwd_dctx(dlogits, ctx.tpi) if (ctx.has_ctx and need[1]) else None
        t_w1, t_b1, t_w2, t_b2, t_wg = ctx.dtypes

        def cast(g, dt):
            return g if g is None or g.dtype == dt else g.to(dt)
        return (dx, dctx, cast(dwg, t_wg), cast(dw1, t_w1), cast(db1, t_b1), cast(dw2, t_w2), cast(db2, t_b2),
                None, None)


def ep_forward(mod, xb, cb, tpi):
    return _EPFunction.apply(xb, cb, mod.gate.weight, mod.w1, mod.b1, mod.w2, mod.b2, mod, tpi)
