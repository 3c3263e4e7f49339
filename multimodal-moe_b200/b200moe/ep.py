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

Overlap: the local experts are cut into `ep_chunks` groups; the send buffer is laid out (group, destination rank,
expert) so that every group has its own contiguous all-to-all.  The all-to-alls run on a side stream, the grouped
GEMMs of group c wait only for group c's rows (CUDA events), so the transfer of group c+1 and the return of group
c-1 overlap the GEMMs of group c.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from . import _lib, ops


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


def chunk_layout(E: int, W: int, chunks: int):
    """Expert order of the send buffer for `chunks` groups of local experts: position of global expert
    e = w*EL + c*ELc + j is c*(W*ELc) + w*ELc + j.  Returns (C, ELc, order[E] int32)."""
    EL = E // W
    C = max(1, min(chunks, EL))
    while EL % C:
        C -= 1
    ELc = EL // C
    e = torch.arange(E)
    w, c, j = e // EL, (e % EL) // ELc, e % ELc
    return C, ELc, (c * (W * ELc) + w * ELc + j).to(torch.int32)


def split_sizes(kept: torch.Tensor, rank: int, W: int, C: int = 1):
    """kept [W,E] (host).  Per chunk c of local experts: rows sent to / received from every rank, and the receive
    segments (source rank major, expert minor).  Returns lists over chunks:
    (send_splits[c][W], recv_splits[c][W], recv_counts[c] int32 [W*ELc], seg_expert int32 [W*ELc])."""
    E = kept.shape[1]
    EL = E // W
    ELc = EL // C
    out_view = kept[rank].view(W, C, ELc)                   # this rank -> (dest rank, chunk, expert)
    in_view = kept[:, rank * EL:(rank + 1) * EL].reshape(W, C, ELc)   # (source rank, chunk, expert) -> this rank
    send = [out_view[:, c].sum(-1).tolist() for c in range(C)]
    recv = [in_view[:, c].sum(-1).tolist() for c in range(C)]
    recv_counts = [in_view[:, c].reshape(-1).to(torch.int32).contiguous() for c in range(C)]
    seg_expert = torch.arange(ELc, dtype=torch.int32).repeat(W)
    return send, recv, recv_counts, seg_expert


def split_sizes_np(kept, rank: int, W: int, C: int = 1):
    """numpy twin of split_sizes for the hot path: kept [W,E] int array -> (send_splits[c][W], recv_splits[c][W])."""
    E = kept.shape[1]
    EL = E // W
    ELc = EL // C
    out_view = kept[rank].reshape(W, C, ELc)
    in_view = kept[:, rank * EL:(rank + 1) * EL].reshape(W, C, ELc)
    send = [out_view[:, c].sum(-1).tolist() for c in range(C)]
    recv = [in_view[:, c].sum(-1).tolist() for c in range(C)]
    return send, recv


def _ep_static(mod, dev, E: int, W: int):
    """Per-module constants of the EP layout, kept on the device: (C, ELc, expert order, segment -> local expert)."""
    key = (str(dev), E, W, getattr(mod, "ep_chunks", 1))
    hit = mod.__dict__.get("_ep_static")
    if hit is None or hit[0] != key:
        C, ELc, order = chunk_layout(E, W, getattr(mod, "ep_chunks", 1))
        seg = torch.arange(ELc, dtype=torch.int32).repeat(W)
        hit = mod.__dict__["_ep_static"] = (key, (C, ELc, order.to(dev), seg.to(dev)))
    return hit[1]


def _pinned(mod, n: int) -> torch.Tensor:
    t = mod.__dict__.get("_ep_pinned")
    if t is None or t.numel() != n:
        t = mod.__dict__["_ep_pinned"] = torch.empty((n,), dtype=torch.int32).pin_memory()
    return t


def _a2a_into(out: torch.Tensor, buf: torch.Tensor, in_splits: List[int], out_splits: List[int], group):
    dist.all_to_all_single(out, buf, output_split_sizes=out_splits, input_split_sizes=in_splits, group=group)


def _comm_stream(mod, dev) -> torch.cuda.Stream:
    s = mod.__dict__.get("_ep_comm_stream")
    if s is None or s.device != dev:
        s = torch.cuda.Stream(device=dev)
        mod.__dict__["_ep_comm_stream"] = s
    return s


def _cumsum0(xs: List[int]) -> List[int]:
    out = [0]
    for v in xs:
        out.append(out[-1] + v)
    return out


class _Chunk:
    """Receive-side state of one group of local experts."""
    __slots__ = ("n_recv", "cap", "offs", "meta", "tile_e", "src", "xp", "pre", "h", "lo", "hi")


class _EPFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x2, ctx_bias, wg, w1, b1, w2, b2, mod, tokens_per_image):
        group = mod.ep_group
        W, rank = dist.get_world_size(group), dist.get_rank(group)
        T, d = x2.shape
        E, k, EL = mod.num_experts, mod.top_k, mod.num_local_experts
        # needs_input_grad is True for parameters even under torch.no_grad(): the caller's grad mode decides
        need_grad = any(ctx.needs_input_grad) and mod.__dict__.get("_grad_mode", True)
        # bf16 operands + gate hi/lo split; the two slab casts run on a side stream under the router and the dispatch
        w1b, w2b, wsplit, wgf, w_ready = mod._prepared(w1, w2, wg, overlap=not _lib.PROFILING)
        b1f, b2f = b1.float().contiguous(), b2.float().contiguous()
        cb = ctx_bias.float().contiguous() if ctx_bias is not None else None
        dev = x2.device
        comp, comm = torch.cuda.current_stream(), _comm_stream(mod, dev)
        st = _lib.stream_ptr()
        i32 = torch.int32

        if w_ready is not None:
            w_ready.fork()
        r = ops.router_fwd(x2, None, cb, tokens_per_image, k, mod.normalize_topk, wg_split=wsplit, E=E)
        if w_ready is not None:
            w_ready.launch()    # after the router: a cast launched first would hold every SM's thread slots
        # histograms of every rank (+ token counts): one kernel, one small all-gather, one kernel for the global
        # routing tables; only the kept counts travel to the host (the one sync of the pass: all-to-all split sizes)
        C, ELc, order_dev, seg_dev = _ep_static(mod, dev, E, W)
        mine = torch.empty((k * E + 1,), dtype=i32, device=dev)
        _lib.call("b200moe_ep_hist", r.block_counts.data_ptr(), r.block_counts.shape[0], k, E, T, mine.data_ptr(), st)
        allh = torch.empty((W, k * E + 1), dtype=i32, device=dev)
        dist.all_gather_into_tensor(allh, mine, group=group)
        goff = torch.empty((k, E), dtype=i32, device=dev)
        kept_dev = torch.empty((W * E + 2,), dtype=i32, device=dev)
        recv_counts_dev = torch.empty((C, W * ELc), dtype=i32, device=dev)
        cf = -1.0 if mod.capacity_factor is None else float(mod.capacity_factor)
        _lib.call("b200moe_ep_route_tables", allh.data_ptr(), W, k, E, rank, cf, ELc, goff.data_ptr(),
                  kept_dev.data_ptr(), recv_counts_dev.data_ptr(), st)
        kept_pin = _pinned(mod, W * E + 2)
        kept_pin.copy_(kept_dev, non_blocking=True)
        done_ev = torch.cuda.Event()
        done_ev.record(comp)
        send_cap = ops.rows_upper_bound(T, k, E, -1)            # host work that does not need the counts
        done_ev.synchronize()
        kept_np = kept_pin.numpy()
        cap = int(kept_np[W * E])
        send_splits, recv_splits = split_sizes_np(kept_np[:W * E].reshape(W, E), rank, W, C)
        send_off = _cumsum0([sum(s_) for s_ in send_splits])
        n_send = send_off[-1]

        p = ops.plan(r, T, k, E, cap, mod.aux_loss_weight, mod.z_loss_weight, global_offset=goff,
                     expert_order=order_dev, row_pad=1, rows_cap=send_cap)
        xs = ops.permute(x2, p.token_of_row, p.meta, send_cap)
        recv_counts = [recv_counts_dev[c] for c in range(C)]

        # dispatch all groups on the communication stream
        ready = torch.cuda.Event()
        ready.record(comp)
        comm.wait_event(ready)
        chunks, arrived = [], []
        with torch.cuda.stream(comm):
            for c in range(C):
                ch = _Chunk()
                ch.n_recv = sum(recv_splits[c])
                ch.lo, ch.hi = c * ELc, (c + 1) * ELc
                xr = torch.empty((max(ch.n_recv, 1), d), dtype=torch.bfloat16, device=dev)
                _a2a_into(xr[:ch.n_recv], xs[send_off[c]:send_off[c + 1]], send_splits[c], recv_splits[c], group)
                ch.xp = xr                                    # replaced by the padded layout below
                ev = torch.cuda.Event()
                ev.record(comm)
                chunks.append(ch)
                arrived.append(ev)
        # grouped GEMMs per group as its rows arrive; results go home on the communication stream
        ys = torch.empty((max(n_send, 1), d), dtype=torch.bfloat16, device=dev)
        if w_ready is not None:
            w_ready.join()
        for c, ch in enumerate(chunks):
            comp.wait_event(arrived[c])
            xr = ch.xp
            xr.record_stream(comp)
            ch.cap = max(128, (ch.n_recv + ELc * 127 + 127) // 128 * 128)
            ch.offs, ch.meta, ch.tile_e, ch.src = ops.ep_recv_plan(recv_counts[c], seg_dev, ELc, ch.cap)
            ch.xp = ops.permute(xr, ch.src, ch.meta, ch.cap)
            ch.pre, ch.h = ops.ffn_up(ch.xp, w1b[ch.lo:ch.hi], b1f[ch.lo:ch.hi], ch.meta, ch.tile_e,
                                      save_pre=need_grad, activation=_lib.ACTIVATIONS[mod.activation])
            yo = ops.ffn_down(ch.h, w2b[ch.lo:ch.hi], b2f[ch.lo:ch.hi], ch.meta, ch.tile_e)
            yr = ops.scatter_rows(yo, ch.src, ch.meta, max(ch.n_recv, 1))
            done = torch.cuda.Event()
            done.record(comp)
            with torch.cuda.stream(comm):
                comm.wait_event(done)
                yr.record_stream(comm)
                _a2a_into(ys[send_off[c]:send_off[c + 1]], yr[:ch.n_recv], recv_splits[c], send_splits[c], group)
        home = torch.cuda.Event()
        home.record(comm)
        comp.wait_event(home)
        y = ops.combine(ys, p.row_of_assignment, r.topk_gate)

        ctx.mod, ctx.tpi, ctx.plan, ctx.routing = mod, tokens_per_image, p, r
        ctx.has_ctx = ctx_bias is not None
        ctx.dtypes = (w1.dtype, b1.dtype, w2.dtype, b2.dtype, wg.dtype)
        ctx.ep = (send_splits, recv_splits, send_off, send_cap, chunks)
        ctx.save_for_backward(x2, wgf, w1b, w2b, ys)
        aux = p.losses[0]
        stats = p.losses.detach()
        # undefined upstream gradients stay None (the engine would otherwise launch a zero fill per missing one, on the
        # critical path between forward and backward); backward handles dy / d_aux == None itself
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(stats)
        mod._stash(r, p)
        return y, aux, stats

    @staticmethod
    def backward(ctx, dy, d_aux, _d_stats):
        x2, wgf, w1b, w2b, ys = ctx.saved_tensors
        mod, p, r = ctx.mod, ctx.plan, ctx.routing
        group = mod.ep_group
        send_splits, recv_splits, send_off, send_cap, chunks = ctx.ep
        n_send = send_off[-1]
        T, d = x2.shape
        k, EL, F = mod.top_k, mod.num_local_experts, mod.dim_feedforward
        need = ctx.needs_input_grad
        dev = x2.device
        comp, comm = torch.cuda.current_stream(), _comm_stream(mod, dev)
        if dy is None:
            dy = torch.zeros_like(x2)
        dy = dy.contiguous().to(torch.bfloat16)
        d_aux_t = d_aux.float().contiguous() if d_aux is not None else None

        dys, dgate_row = ops.combine_bwd(dy, ys, p.token_of_row, p.gate_of_row, p.meta, send_cap)
        ready = torch.cuda.Event()
        ready.record(comp)
        comm.wait_event(ready)
        arrived, dyrs = [], []
        with torch.cuda.stream(comm):
            for c, ch in enumerate(chunks):
                dyr = torch.empty((max(ch.n_recv, 1), d), dtype=torch.bfloat16, device=dev)
                _a2a_into(dyr[:ch.n_recv], dys[send_off[c]:send_off[c + 1]], send_splits[c], recv_splits[c], group)
                ev = torch.cuda.Event()
                ev.record(comm)
                arrived.append(ev)
                dyrs.append(dyr)
        # router gradient does not depend on the experts: it fills the bubble while the first rows travel
        dlogits = ops.router_bwd_dlogits(r.logits, r.topk_idx, p.row_of_assignment, dgate_row, p.tokens_per_expert,
                                         d_aux_t, mod.normalize_topk, mod.aux_loss_weight, mod.z_loss_weight)
        dwg, dctx = ops.router_bwd_params(dlogits, x2, ctx.tpi, want_dwg=need[2],
                                          want_dctx=bool(ctx.has_ctx and need[1]))

        f32 = torch.float32
        dw1 = torch.empty((EL, F, d), dtype=f32, device=dev) if need[3] else None
        db1 = torch.empty((EL, F), dtype=f32, device=dev) if need[4] else None
        dw2 = torch.empty((EL, d, F), dtype=f32, device=dev) if need[5] else None
        db2 = torch.empty((EL, d), dtype=f32, device=dev) if need[6] else None
        dxs = torch.empty((max(n_send, 1), d), dtype=torch.bfloat16, device=dev) if need[0] else None
        for c, ch in enumerate(chunks):
            comp.wait_event(arrived[c])
            dyrs[c].record_stream(comp)
            n = ch.hi - ch.lo
            dyp = ops.permute(dyrs[c], ch.src, ch.meta, ch.cap)
            dhpre = ops.ffn_dgrad_down(dyp, w2b[ch.lo:ch.hi], ch.pre, ch.meta, ch.tile_e,
                                       activation=_lib.ACTIVATIONS[mod.activation])
            if need[0]:
                dxp = ops.ffn_dgrad_up(dhpre, w1b[ch.lo:ch.hi], ch.meta, ch.tile_e)
                dxr = ops.scatter_rows(dxp, ch.src, ch.meta, max(ch.n_recv, 1))
                done = torch.cuda.Event()
                done.record(comp)
                with torch.cuda.stream(comm):
                    comm.wait_event(done)
                    dxr.record_stream(comm)
                    _a2a_into(dxs[send_off[c]:send_off[c + 1]], dxr[:ch.n_recv], recv_splits[c], send_splits[c],
                              group)
            if need[5] and need[3]:
                ops.ffn_wgrad2(dyp, ch.h, dhpre, ch.xp, ch.offs, n, out0=dw2[ch.lo:ch.hi], out1=dw1[ch.lo:ch.hi])
            elif need[5]:
                ops.ffn_wgrad(dyp, ch.h, ch.offs, n, out=dw2[ch.lo:ch.hi])
            elif need[3]:
                ops.ffn_wgrad(dhpre, ch.xp, ch.offs, n, out=dw1[ch.lo:ch.hi])
            if need[6] and need[4]:
                ops.colsum(dyp, ch.offs, n, out=db2[ch.lo:ch.hi], a1=dhpre, out1=db1[ch.lo:ch.hi])
            elif need[6]:
                ops.colsum(dyp, ch.offs, n, out=db2[ch.lo:ch.hi])
            elif need[4]:
                ops.colsum(dhpre, ch.offs, n, out=db1[ch.lo:ch.hi])
        dx = None
        if need[0]:
            home = torch.cuda.Event()
            home.record(comm)
            comp.wait_event(home)
            dx = ops.unpermute_bwd(dxs, p.row_of_assignment, dlogits, wgf, T, k, d)
        t_w1, t_b1, t_w2, t_b2, t_wg = ctx.dtypes

        scale = float(getattr(mod, "expert_grad_scale", 1.0))

        def cast(g, dt, expert=False):
            if g is not None and expert and scale != 1.0:
                g = g.mul_(scale)                 # DDP's mean-over-ranks convention (b200moe/checkpoint.py)
            return g if g is None or g.dtype == dt else g.to(dt)
        return (dx, dctx, cast(dwg, t_wg), cast(dw1, t_w1, True), cast(db1, t_b1, True), cast(dw2, t_w2, True),
                cast(db2, t_b2, True),
                None, None)


def ep_forward(mod, xb, cb, tpi):
    return _EPFunction.apply(xb, cb, mod.gate.weight, mod.w1, mod.b1, mod.w2, mod.b2, mod, tpi)
