"""Expert parallelism over peer memory: the row-movement kernels of the layer load and store rows in
the other ranks' HBM over NVLink, so there is no all-to-all, no send / receive buffer and no host synchronisation in
the pass (csrc/ep_p2p.cu has the data flow; b200moe/ep.py is the NCCL all-to-all path this replaces on one NVSwitch box
and remains the fallback).

Per module one symmetric arena (torch.distributed._symmetric_memory: cuMem allocations mapped into every rank of the
group, plus signal pads for a device-side barrier) holds xp | yo | dyp | dxp, each [rows_cap, d] bf16 with
rows_cap = E_local * pad(capacity for T_cap tokens per rank).  T_cap is agreed once: the constructor's `ep_max_tokens`
(declare it for multiscale training) or else the first batch, max over the ranks.  What has to fit is the GLOBAL row
count, so single ranks may exceed T_cap; if the rows of a pass do not fit, the device-side flag of ep_p2p_tables (the
same on every rank) makes the next backward / forward raise on ALL ranks.

Forward : router -> histogram all-gather -> ep_route_tables -> plan (send layout) -> ep_p2p_tables ->
          ep_p2p_dispatch (gather x rows, store them into the expert ranks' xp) -> barrier -> up -> down (into yo)
          -> barrier -> ep_p2p_combine (load the k rows of a token from the expert ranks' yo).
Backward: ep_p2p_combine_bwd (store gate * dy into the expert ranks' dyp; load yo rows for the gate gradient) ->
          router gradients (fill the wait) -> barrier -> dgrad_down -> dgrad_up (into dxp) ->
          barrier -> ep_p2p_pull_rows on a side stream (copy the dxp rows home over NVLink) next to the weight
          gradients on the main stream -> unpermute_bwd.
Reuse of the arena across steps is ordered by the histogram all-gather of the next forward (every rank has finished
the previous backward before any rank's all-gather completes).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import _lib, ops

_BIG = 1 << 30


def _side_stream(mod, dev, which: int = 0) -> torch.cuda.Stream:
    key = "_p2p_side" if which == 0 else f"_p2p_side{which}"
    s = mod.__dict__.get(key)
    if s is None or s.device != dev:
        s = mod.__dict__[key] = torch.cuda.Stream(device=dev)
    return s


def p2p_layout(kept: torch.Tensor, rank: int, row_pad: int = 128):
    """Host twin of b200moe_ep_p2p_tables (pure integer logic, pinned on CPU with gloo and against the kernel on GPU).
    kept [W,E]: rows every rank sends to every expert.  Returns (dest_off [E]: first row, inside the buffer of rank
    e // E_local, of `rank`'s rows for expert e; expert_offsets [W, E_local+1]: the padded segments of every rank)."""
    W, E = kept.shape
    EL = E // W
    kept = kept.to(torch.int64)
    tot = kept.sum(0)
    seg = ((tot + row_pad - 1) // row_pad * row_pad).view(W, EL)
    start = seg.cumsum(1) - seg                                  # [W,EL] first row of a local expert on its rank
    before = kept.cumsum(0) - kept                               # [W,E] rows of the lower ranks (rank-major in an expert)
    dest_off = start.reshape(E) + before[rank]
    offsets = torch.cat([start, seg.sum(1, keepdim=True)], 1)
    return dest_off.to(torch.int32), offsets.to(torch.int32)


class P2PUnavailable(RuntimeError):
    pass


def _agree(ok: bool, group, dev) -> bool:
    """True iff `ok` on every rank (every rank calls this at the same point, whatever happened to it locally)."""
    t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    return int(t.item()) == 1


class _Arena:
    """Symmetric row buffers of one module.  Set-up is a fixed sequence of collectives that every rank executes in the
    same order whether or not a local step failed, so a failure on one rank turns into P2PUnavailable on all of them
    instead of a hang in mismatched collectives."""

    def __init__(self, mod, group, T_local: int, dev):
        err, symm = None, None
        try:
            import torch.distributed._symmetric_memory as symm      # private torch API: probed, never assumed
        except Exception as e:  # noqa: BLE001
            err = e
        if not _agree(err is None, group, dev):
            raise P2PUnavailable(f"torch symmetric memory is not importable on every rank: {err!r}")
        W = dist.get_world_size(group)
        E, k, d = mod.num_experts, mod.top_k, mod.d_model
        # tokens per rank the arena is sized for: the declared maximum (multiscale training varies the token count per
        # step and per rank, configs/rtdetrv2/include/dataloader.yml:22) or else the first batch, max over the ranks
        t = torch.tensor([max(int(T_local), int(mod.ep_max_tokens or 0))], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        self.T_cap = int(t.item())
        cap = ops.capacity_of(mod.capacity_factor, self.T_cap * W, k, E)
        # 256-row segments (CTA-pair GEMM tiles) when a balanced share of the rows still fills two waves of them
        self.row_pad = int(_lib.load().b200moe_layer_row_pad(self.T_cap, k, mod.num_local_experts, -1, d,
                                                             mod.dim_feedforward))
        self.rows_cap = mod.num_local_experts * ((cap + self.row_pad - 1) // self.row_pad * self.row_pad)
        self.buf_bytes = self.rows_cap * d * 2
        try:
            self.mem = symm.empty(4 * self.buf_bytes, dtype=torch.uint8, device=dev)
        except Exception as e:  # noqa: BLE001
            err = e
        if not _agree(err is None, group, dev):
            raise P2PUnavailable(f"symmetric allocation of {4 * self.buf_bytes} bytes failed on some rank: {err!r}")
        try:
            self.hdl = symm.rendezvous(self.mem, group)               # collective: reached by every rank or by none
            self.peers = torch.tensor(list(self.hdl.buffer_ptrs), dtype=torch.int64, device=dev)
        except Exception as e:  # noqa: BLE001
            err = e
        if not _agree(err is None, group, dev):
            raise P2PUnavailable(f"symmetric-memory rendezvous failed on some rank: {err!r}")
        bf = self.mem.view(torch.bfloat16).view(4, self.rows_cap, d)
        self.xp, self.yo, self.dyp, self.dxp = bf[0], bf[1], bf[2], bf[3]
        self.off = [i * self.buf_bytes for i in range(4)]
        self.pending = False           # a forward whose backward has not run yet owns the arena
        # overflow evidence: meta[5] of b200moe_ep_p2p_tables (identical on every rank, it is computed from the gathered
        # histogram) lands here asynchronously and is looked at by the next backward / forward
        self.flag_host = torch.zeros(8, dtype=torch.int32).pin_memory()
        self.flag_event = None

    def barrier(self, channel: int):
        self.hdl.barrier(channel=channel)

    def note_meta(self, meta: torch.Tensor):
        self.flag_host.copy_(meta, non_blocking=True)
        self.flag_event = torch.cuda.Event()
        self.flag_event.record()

    def check_overflow(self):
        """Raises on EVERY rank (the flag is the same everywhere) when the rows of the last pass did not fit the arena."""
        if self.flag_event is None:
            return
        self.flag_event.synchronize()
        self.flag_event = None
        if int(self.flag_host[5]) != 0:
            raise RuntimeError(
                f"expert-parallel peer-memory arena overflow: it was sized for {self.T_cap} tokens per rank "
                f"({self.rows_cap} rows), the last batch needed more.  Construct the layer with ep_max_tokens=<largest "
                f"tokens per rank> (multiscale training: the largest scale), or with ep_p2p=False for the NCCL path")


def arena_for(mod, T: int, dev) -> _Arena:
    a = mod.__dict__.get("_p2p_arena")
    if a is None:
        a = mod.__dict__["_p2p_arena"] = _Arena(mod, mod.ep_group, T, dev)    # raises P2PUnavailable on every rank or none
    # no per-rank check of T against T_cap here: a rank-local raise would leave the other ranks waiting in the
    # histogram all-gather.  What must fit is the GLOBAL row count; the device-side flag says so on all ranks alike.
    a.check_overflow()
    return a


class _P2PFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x2, ctx_bias, wg, w1, b1, w2, b2, mod, tokens_per_image, ar):
        group = mod.ep_group
        W, rank = dist.get_world_size(group), dist.get_rank(group)
        T, d = x2.shape
        E, k, EL = mod.num_experts, mod.top_k, mod.num_local_experts
        # needs_input_grad is True for parameters even under torch.no_grad(): the caller's grad mode decides
        need_grad = any(ctx.needs_input_grad) and mod.__dict__.get("_grad_mode", True)
        # the bf16 casts of this rank's expert slabs run on a side stream under the router, the histogram exchange and
        # the dispatch; the up-projection is the first kernel that needs them
        w1b, w2b, wsplit, wgf, ready = mod._prepared(w1, w2, wg, overlap=not _lib.PROFILING)
        b1f, b2f = b1.float().contiguous(), b2.float().contiguous()
        cb = ctx_bias.float().contiguous() if ctx_bias is not None else None
        dev = x2.device
        st = _lib.stream_ptr()
        i32 = torch.int32

        if ready is not None:
            ready.fork()
        r = ops.router_fwd(x2, None, cb, tokens_per_image, k, mod.normalize_topk, wg_split=wsplit, E=E)
        if ready is not None:
            ready.launch()      # after the router: a cast launched first would hold every SM's thread slots
        mine = torch.empty((k * E + 1,), dtype=i32, device=dev)
        _lib.call("b200moe_ep_hist", r.block_counts.data_ptr(), r.block_counts.shape[0], k, E, T, mine.data_ptr(), st)
        allh = torch.empty((W, k * E + 1), dtype=i32, device=dev)
        dist.all_gather_into_tensor(allh, mine, group=group)
        goff = torch.empty((k, E), dtype=i32, device=dev)
        kept = torch.empty((W * E + 2,), dtype=i32, device=dev)
        recv_counts = torch.empty((W * EL,), dtype=i32, device=dev)
        cf = -1.0 if mod.capacity_factor is None else float(mod.capacity_factor)
        _lib.call("b200moe_ep_route_tables", allh.data_ptr(), W, k, E, rank, cf, EL, goff.data_ptr(), kept.data_ptr(),
                  recv_counts.data_ptr(), st)
        # the plan kernel takes the capacity from the host; the global capacity lives on the device (kept[W*E]), so the
        # offsets are shifted instead: room = BIG - (offset + BIG - capacity) = capacity - offset
        goff_shift = goff + (_BIG - kept[W * E])
        send_cap = ops.rows_upper_bound(T, k, E, -1)
        p = ops.plan(r, T, k, E, _BIG, mod.aux_loss_weight, mod.z_loss_weight, global_offset=goff_shift,
                     expert_order=None, row_pad=1, rows_cap=send_cap)
        p.capacity = kept[W * E]          # the global capacity (device scalar; the plan call above only saw the shifted offsets)
        R = ar.rows_cap
        dest_off = torch.empty((E,), dtype=i32, device=dev)
        offs = torch.empty((EL + 1,), dtype=i32, device=dev)
        meta = torch.empty((8,), dtype=i32, device=dev)
        tile_e = torch.empty((R // 128,), dtype=i32, device=dev)
        _lib.call("b200moe_ep_p2p_tables", kept.data_ptr(), W, E, rank, R, ar.row_pad, dest_off.data_ptr(), offs.data_ptr(),
                  meta.data_ptr(), tile_e.data_ptr(), st)
        ar.note_meta(meta)
        _lib.call("b200moe_ep_p2p_zero_pads", kept.data_ptr(), offs.data_ptr(), W, E, rank, d, ar.xp.data_ptr(),
                  ar.dyp.data_ptr() if need_grad else None, st)
        dest_of_row = torch.empty((send_cap,), dtype=i32, device=dev)
        _lib.call("b200moe_ep_p2p_dispatch", x2.data_ptr(), p.token_of_row.data_ptr(), p.expert_offsets.data_ptr(),
                  p.meta.data_ptr(), dest_off.data_ptr(), ar.peers.data_ptr(), ar.off[0], R, send_cap, E, EL, d,
                  dest_of_row.data_ptr(), st)
        ar.barrier(0)                                   # every rank's rows have landed in xp
        if ready is not None:
            ready.join()
        pre, h = ops.ffn_up(ar.xp, w1b, b1f, meta, tile_e, save_pre=need_grad, row_pad=ar.row_pad,
                            activation=_lib.ACTIVATIONS[mod.activation])
        ops.ffn_down(h, w2b, b2f, meta, tile_e, row_pad=ar.row_pad, out=ar.yo)
        ar.barrier(1)                                   # every rank's yo is complete
        y = torch.empty((T, d), dtype=torch.bfloat16, device=dev)
        _lib.call("b200moe_ep_p2p_combine", ar.peers.data_ptr(), ar.off[1], R, dest_of_row.data_ptr(),
                  p.row_of_assignment.data_ptr(), r.topk_gate.data_ptr(), T, k, d, y.data_ptr(), st)

        ctx.mod, ctx.tpi, ctx.plan, ctx.routing, ctx.ar = mod, tokens_per_image, p, r, ar
        ctx.has_ctx = ctx_bias is not None
        ctx.dtypes = (w1.dtype, b1.dtype, w2.dtype, b2.dtype, wg.dtype)
        ctx.tabs = (dest_of_row, offs, meta, tile_e, send_cap, pre, h)
        ctx.save_for_backward(x2, wgf, w1b, w2b)
        ar.pending = need_grad
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
        x2, wgf, w1b, w2b = ctx.saved_tensors
        mod, p, r, ar = ctx.mod, ctx.plan, ctx.routing, ctx.ar
        dest_of_row, offs, meta, tile_e, send_cap, pre, h = ctx.tabs
        T, d = x2.shape
        E, k, EL, F = mod.num_experts, mod.top_k, mod.num_local_experts, mod.dim_feedforward
        need = ctx.needs_input_grad
        dev = x2.device
        st = _lib.stream_ptr()
        R = ar.rows_cap
        if dy is None:
            dy = torch.zeros_like(x2)
        dy = dy.contiguous().to(torch.bfloat16)
        d_aux_t = d_aux.float().contiguous() if d_aux is not None else None
        f32 = torch.float32

        ar.check_overflow()               # before any gradient of an overflowed pass reaches the optimizer
        dgate_row = torch.empty((send_cap,), dtype=f32, device=dev)
        _lib.call("b200moe_ep_p2p_combine_bwd", dy.data_ptr(), ar.peers.data_ptr(), ar.off[1], ar.off[2], R,
                  dest_of_row.data_ptr(), p.token_of_row.data_ptr(), p.gate_of_row.data_ptr(), p.meta.data_ptr(),
                  send_cap, d, dgate_row.data_ptr(), st)
        # dlogits feeds unpermute_bwd at the end of the dx chain; the parameter gradients of the router (the gate's weight
        # gradient is 2 T E d fp32 FMAs on CUDA cores: 106 us at workload D) wait for the tail of the pass, where they run
        # next to the bandwidth-bound bias sums and unpermute_bwd instead of in front of the barrier (round-2 timeline)
        dlogits = ops.router_bwd_dlogits(r.logits, r.topk_idx, p.row_of_assignment, dgate_row, p.tokens_per_expert,
                                         d_aux_t, mod.normalize_topk, mod.aux_loss_weight, mod.z_loss_weight)
        ar.barrier(2)                                   # every rank's dyp rows have landed
        dhpre = ops.ffn_dgrad_down(ar.dyp, w2b, pre, meta, tile_e, row_pad=ar.row_pad,
                                   activation=_lib.ACTIVATIONS[mod.activation])
        comp = torch.cuda.current_stream()
        overlap = not _lib.PROFILING                    # per-kernel timing wants kernels running alone
        side = _side_stream(mod, dev) if overlap else comp          # rows home, then the bias sums
        side2 = _side_stream(mod, dev, 1) if overlap else comp      # router parameter gradients
        dxs = None
        if need[0]:
            ops.ffn_dgrad_up(dhpre, w1b, meta, tile_e, row_pad=ar.row_pad, out=ar.dxp)
            ar.barrier(3)                               # every rank's dxp is complete
        side.wait_stream(comp)
        pulled = None
        if need[0]:
            # the rows come home on a side stream (a plain copy kernel: no shared memory, so it runs on the SMs next to
            # the persistent weight-gradient CTAs) while the tensor-bound kernels below keep the main stream busy
            dxs = torch.empty((send_cap, d), dtype=torch.bfloat16, device=dev)
            with torch.cuda.stream(side):
                _lib.call("b200moe_ep_p2p_pull_rows", ar.peers.data_ptr(), ar.off[3], R, dest_of_row.data_ptr(),
                          p.meta.data_ptr(), send_cap, d, dxs.data_ptr(), side.cuda_stream)
                pulled = torch.cuda.Event()
                pulled.record(side)
        dw1 = torch.empty((EL, F, d), dtype=f32, device=dev) if need[3] else None
        db1 = torch.empty((EL, F), dtype=f32, device=dev) if need[4] else None
        dw2 = torch.empty((EL, d, F), dtype=f32, device=dev) if need[5] else None
        db2 = torch.empty((EL, d), dtype=f32, device=dev) if need[6] else None
        if need[5] and need[3]:
            ops.ffn_wgrad2(ar.dyp, h, dhpre, ar.xp, offs, EL, out0=dw2, out1=dw1)     # one persistent grid for both
        elif need[5]:
            ops.ffn_wgrad(ar.dyp, h, offs, EL, out=dw2)
        elif need[3]:
            ops.ffn_wgrad(dhpre, ar.xp, offs, EL, out=dw1)
        if need[4] or need[6]:
            # bias gradients: a bandwidth-bound column sum behind the returning rows on the side stream
            with torch.cuda.stream(side):
                if need[6] and need[4]:
                    ops.colsum(ar.dyp, offs, EL, out=db2, a1=dhpre, out1=db1)
                elif need[6]:
                    ops.colsum(ar.dyp, offs, EL, out=db2)
                else:
                    ops.colsum(dhpre, offs, EL, out=db1)
        dwg = dctx = None
        if need[2] or (ctx.has_ctx and need[1]):
            side2.wait_stream(comp)                     # behind the weight gradients: the tail of the pass
            with torch.cuda.stream(side2):
                dwg, dctx = ops.router_bwd_params(dlogits, x2, ctx.tpi, want_dwg=need[2],
                                                  want_dctx=bool(ctx.has_ctx and need[1]))
        dx = None
        if need[0]:
            if pulled is not None and overlap:
                comp.wait_event(pulled)                 # only the rows: the bias sums keep running beside unpermute_bwd
            dx = ops.unpermute_bwd(dxs, p.row_of_assignment, dlogits, wgf, T, k, d)
        if overlap:
            comp.wait_stream(side)
            comp.wait_stream(side2)
        ar.pending = False
        t_w1, t_b1, t_w2, t_b2, t_wg = ctx.dtypes

        scale = float(getattr(mod, "expert_grad_scale", 1.0))

        def cast(g, dt, expert=False):
            if g is not None and expert and scale != 1.0:
                g = g.mul_(scale)                 # DDP's mean-over-ranks convention (b200moe/checkpoint.py)
            return g if g is None or g.dtype == dt else g.to(dt)
        return (dx, dctx, cast(dwg, t_wg), cast(dw1, t_w1, True), cast(db1, t_b1, True), cast(dw2, t_w2, True),
                cast(db2, t_b2, True),
                None, None, None)


def p2p_forward(mod, xb, cb, tpi, ar):
    return _P2PFunction.apply(xb, cb, mod.gate.weight, mod.w1, mod.b1, mod.w2, mod.b2, mod, tpi, ar)
