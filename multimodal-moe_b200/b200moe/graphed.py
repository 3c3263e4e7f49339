"""CUDA-graph replay of a ContextMoE for fixed-shape training (RT-DETR trains at one resolution unless multi-scale is
on: configs/rtdetrv2/include/dataloader.yml:22, so the S5 encoder level sees the same [B, HW, d] every step).

At the reference's shapes (8 x 400 tokens, d = 256) one forward + backward of the layer is ~0.13 ms of device work
behind ~0.5 ms of Python / autograd / allocator time when launched call by call.  `GraphedMoE` captures the forward
(weight preparation, context bias, b200moe_layer_fwd) and the backward (b200moe_layer_bwd, context-bias gradients)
once each into CUDA graphs that share one memory pool, and replays them from a thin autograd node: per step the host
copies the inputs into the captured buffers, launches one graph each way and hands out the captured outputs.

    moe = ContextMoE(...).cuda()
    fast = GraphedMoE(moe, x_example, context_example)      # captures; x_example gives shape and dtype
    y = fast(x, context); loss = crit(y) + fast.aux_loss; loss.backward()      # as with the module itself

Semantics are those of `torch.cuda.make_graphed_callables`: the returned tensors (and the parameter gradients autograd
accumulates from) are the captured buffers, valid until the next replay; shapes, dtypes, `requires_grad` pattern and the
parameter storage must not change (in-place optimizer steps are fine; weights are re-cast inside the graph every step).
Anything else — other shapes, no-grad evaluation, expert parallelism (its pass has a host sync) — falls through to the
eager module.  Parity with the eager module is bit-exact (tests/test_gpu_layer.py::test_graphed_matches_eager).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from .layer import ContextMoE


class _Replay(torch.autograd.Function):
    @staticmethod
    def forward(ctx, g, x, context, *params):
        g.x_s.copy_(x)
        if context is not None:
            g.ctx_s.copy_(context)
        g.fwd_graph.replay()
        ctx.g = g
        y, aux = g.y_s.detach(), g.aux_s.detach()
        return y, aux

    @staticmethod
    def backward(ctx, dy, daux):
        g = ctx.g
        if dy is None:
            g.dy_s.zero_()
        else:
            g.dy_s.copy_(dy)
        if daux is None:
            g.daux_s.zero_()
        else:
            g.daux_s.copy_(daux)
        for p, s in zip(g._params, g.dparams_s):
            # autograd adopts a returned buffer as p.grad when it can; if a gradient from an earlier step still lives
            # there (accumulation over micro-batches, zero_grad(set_to_none=False)), move it out before the replay
            if s is not None and p.grad is not None and p.grad.data_ptr() == s.data_ptr():
                p.grad = p.grad.clone()
        g.bwd_graph.replay()
        return (None, g.dx_s.detach() if g.dx_s is not None else None, None,
                *[p.detach() if p is not None else None for p in g.dparams_s])


class GraphedMoE(nn.Module):
    def __init__(self, moe: ContextMoE, x_example: torch.Tensor, context_example: Optional[torch.Tensor] = None,
                 warmup: int = 3):
        super().__init__()
        if not isinstance(moe, ContextMoE):
            raise TypeError("GraphedMoE wraps a b200moe.ContextMoE")
        if moe.ep_size > 1:
            raise ValueError("the expert-parallel pass synchronises with the host once per step and cannot be captured")
        if not x_example.is_cuda:
            raise RuntimeError("b200moe runs on a B200 (sm_100a) only; there is no CPU fallback")
        self.moe = moe
        self._shape, self._dtype = tuple(x_example.shape), x_example.dtype
        self._has_ctx = context_example is not None
        self._params = [p for p in moe.parameters() if p.requires_grad]
        self._capture(x_example, context_example, warmup)

    @property
    def aux_loss(self) -> torch.Tensor:
        return self.moe.aux_loss

    @property
    def last_aux(self):
        return self.moe.last_aux

    def _capture(self, x_example, context_example, warmup):
        moe, params = self.moe, self._params
        dev = x_example.device
        self.x_s = torch.zeros_like(x_example).requires_grad_(True)
        with torch.no_grad():
            self.x_s.copy_(x_example)
        self.ctx_s = context_example.clone() if self._has_ctx else None
        self.dy_s = torch.zeros_like(x_example)
        self.daux_s = torch.ones((), dtype=torch.float32, device=dev)
        inputs = (self.x_s, *params)

        def fwd():
            moe._wcache.clear()                    # the fp32 -> bf16 weight preparation belongs to the graph
            y = moe(self.x_s, self.ctx_s)
            return y, moe.aux_loss

        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):
                y, aux = fwd()
                torch.autograd.grad((y, aux), inputs, (self.dy_s, self.daux_s), allow_unused=True)
            del y, aux
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)

        pool = torch.cuda.graph_pool_handle()
        self.fwd_graph, self.bwd_graph = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.fwd_graph, pool=pool):
            y, aux = fwd()
        with torch.cuda.graph(self.bwd_graph, pool=pool):
            grads = torch.autograd.grad((y, aux), inputs, (self.dy_s, self.daux_s), allow_unused=True)
        self.y_s, self.aux_s = y, aux
        self.dx_s, self.dparams_s = grads[0], list(grads[1:])
        self._telemetry = moe.last_aux             # views into the captured workspace: refreshed by every replay
        torch.cuda.synchronize(dev)

    def forward(self, x: torch.Tensor, context: Optional[torch.Tensor] = None) -> torch.Tensor:
        moe = self.moe
        if (not torch.is_grad_enabled() or tuple(x.shape) != self._shape or x.dtype != self._dtype
                or (context is not None) != self._has_ctx
                or (context is not None and context.shape != self.ctx_s.shape)):
            return moe(x, context)
        y, aux = _Replay.apply(self, x, context, *self._params)
        moe.aux_loss, moe.last_aux = aux, self._telemetry
        return y


class GraphedStep:
    """Whole training step of the layer -- forward, the caller's loss, backward -- as ONE CUDA graph (PyTorch's
    "whole-network capture" recipe packaged for this layer).  Where `GraphedMoE` still crosses the autograd engine twice
    per step (~0.2 ms of host time at the reference's 8 x 400 tokens against ~0.1 ms of device time), this replays the
    step with a single launch:

        step = GraphedStep(moe, loss_fn, x_example, context_example, *extra_examples)
        loss = step(x, context, *extra)          # x / context / extra may be pinned HOST tensors: they are copied
        opt.step(); ...                          #   straight into the captured inputs (one H2D each, no staging)

    `loss_fn(y, *extra) -> scalar` is the caller's criterion on the layer output; the weighted auxiliary loss of the
    layer is added to it (det_engine.py:51 sums the loss dict).  After the call every parameter's `.grad` holds the
    gradient of this step (static buffers: written, not accumulated, by each replay), `step.dx` the gradient with respect
    to x, and the returned 0-dim tensor is the captured loss (read it with .item()).  Shapes, dtypes and the parameter
    storage must not change; in-place optimizer steps are fine (weights are re-cast inside the graph).
    """

    def __init__(self, moe: ContextMoE, loss_fn, x_example: torch.Tensor, context_example: Optional[torch.Tensor] = None,
                 *extra_examples: torch.Tensor, warmup: int = 3):
        if not isinstance(moe, ContextMoE):
            raise TypeError("GraphedStep wraps a b200moe.ContextMoE")
        if moe.ep_size > 1:
            raise ValueError("the expert-parallel pass cannot be captured (its transport set-up synchronises with the host)")
        dev = next(moe.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("b200moe runs on a B200 (sm_100a) only; there is no CPU fallback")
        self.moe, self._loss_fn = moe, loss_fn
        self.x_s = x_example.to(dev).clone().requires_grad_(True)
        self.ctx_s = context_example.to(dev).clone() if context_example is not None else None
        self.extra_s = [e.to(dev).clone() for e in extra_examples]
        params = [p for p in moe.parameters() if p.requires_grad]

        def run():
            moe._wcache.clear()                    # the fp32 -> bf16 weight preparation belongs to the graph
            y = moe(self.x_s, self.ctx_s)
            loss = self._loss_fn(y, *self.extra_s) + moe.aux_loss
            loss.backward()
            return loss.detach()

        for p in params:
            p.grad = None
        self.x_s.grad = None
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):
                run()
                for p in params:
                    p.grad = None
                self.x_s.grad = None
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss_s = run()                    # .grad tensors are allocated from the graph's pool: static from here on
        self._params, self._grads = params, [p.grad for p in params]
        self._dx = self.x_s.grad
        self._telemetry = moe.last_aux
        torch.cuda.synchronize(dev)

    @property
    def dx(self) -> Optional[torch.Tensor]:
        return self._dx

    def __call__(self, x: torch.Tensor, context: Optional[torch.Tensor] = None, *extra: torch.Tensor) -> torch.Tensor:
        with torch.no_grad():
            self.x_s.copy_(x, non_blocking=True)
            if self.ctx_s is not None and context is not None:
                self.ctx_s.copy_(context, non_blocking=True)
            for s, e in zip(self.extra_s, extra):
                s.copy_(e, non_blocking=True)
        return self.replay()

    def replay(self) -> torch.Tensor:
        """Replays the step on whatever the captured input buffers (`x_s`, `ctx_s`, `extra_s`) hold."""
        self.graph.replay()
        for p, g in zip(self._params, self._grads):       # optimizer.zero_grad(set_to_none=True) detaches them
            if p.grad is not g:
                p.grad = g
        self.moe.last_aux = self._telemetry
        return self.loss_s


class PipelinedGraphedStep:
    """`GraphedStep` with the input copy of batch i+1 overlapped with the step on batch i: two captured steps with their
    own input buffers, a copy stream, and events in both directions (a buffer is refilled only after the replay that read
    it, a replay starts only after its buffer is filled).

        step = PipelinedGraphedStep(moe, loss_fn, x_example, context_example, *extra_examples)
        step.prefetch(x0, c0, *e0)
        for batch in loader:                       # batch = the NEXT inputs (pinned host tensors)
            loss = step()                          # launches the step on the batch prefetched last
            step.prefetch(*batch)                  # its copies run on the copy stream under that step
            opt.step(); print(loss.item())

    Gradients land in `.grad` as with `GraphedStep` (each of the two captured steps owns a set of gradient buffers; the
    one of the step just replayed is attached)."""

    def __init__(self, moe: ContextMoE, loss_fn, x_example: torch.Tensor, context_example: Optional[torch.Tensor] = None,
                 *extra_examples: torch.Tensor, warmup: int = 3):
        self._steps = [GraphedStep(moe, loss_fn, x_example, context_example, *extra_examples, warmup=warmup)
                       for _ in range(2)]
        dev = self._steps[0].x_s.device
        self._copy = torch.cuda.Stream(device=dev)
        self._ready = [torch.cuda.Event() for _ in range(2)]
        self._done = [torch.cuda.Event() for _ in range(2)]
        for e in self._done:
            e.record(torch.cuda.current_stream(dev))
        self._next, self._pending = 0, None
        self.moe = moe

    @property
    def dx(self) -> Optional[torch.Tensor]:
        return self._steps[1 - self._next].dx             # of the step replayed last

    def prefetch(self, x: torch.Tensor, context: Optional[torch.Tensor] = None, *extra: torch.Tensor) -> None:
        if self._pending is not None:
            raise RuntimeError("PipelinedGraphedStep.prefetch: the batch prefetched before has not been run yet")
        j = self._next
        s = self._steps[j]
        self._copy.wait_event(self._done[j])              # the last replay that read these buffers
        with torch.cuda.stream(self._copy), torch.no_grad():
            s.x_s.copy_(x, non_blocking=True)
            if s.ctx_s is not None and context is not None:
                s.ctx_s.copy_(context, non_blocking=True)
            for b, e in zip(s.extra_s, extra):
                b.copy_(e, non_blocking=True)
            self._ready[j].record(self._copy)
        self._pending = j

    def __call__(self) -> torch.Tensor:
        if self._pending is None:
            raise RuntimeError("PipelinedGraphedStep: call prefetch(x, context, ...) before the step")
        j, self._pending = self._pending, None
        s = self._steps[j]
        cur = torch.cuda.current_stream(s.x_s.device)
        cur.wait_event(self._ready[j])
        loss = s.replay()
        self._done[j].record(cur)
        self._next = 1 - j
        return loss


class MultiShapeGraphedMoE(nn.Module):
    """`GraphedMoE` for training loops whose token count changes from step to step.  The vendored recipe trains
    multi-scale (configs/rtdetrv2/include/dataloader.yml:22: a scale out of 480 .. 800 in steps of 32 per batch, i.e. about
    a dozen distinct [B, HW, d] at the S5 level), which a single captured shape cannot serve; launched call by call the
    layer costs ~0.5 ms of host time per step against ~0.1 ms on the device.  This wrapper keeps one `GraphedMoE` per
    (shape, dtype, context shape), captured the first time the shape is seen, and replays it afterwards:

        fast = MultiShapeGraphedMoE(moe, max_shapes=16)
        y = fast(x, context); loss = crit(y) + fast.aux_loss; loss.backward()

    Every cached shape owns its captured workspaces and gradient buffers (about the size of the layer's parameters in
    fp32 plus the activations of that shape), hence `max_shapes`; shapes beyond it, no-grad calls and expert-parallel
    layers run the eager module.  Results are those of the eager module bit for bit
    (tests/test_gpu_layer.py::test_multishape_graphed_matches_eager)."""

    def __init__(self, moe: ContextMoE, max_shapes: int = 16, warmup: int = 3):
        super().__init__()
        if not isinstance(moe, ContextMoE):
            raise TypeError("MultiShapeGraphedMoE wraps a b200moe.ContextMoE")
        self.moe = moe
        self.max_shapes, self._warmup = int(max_shapes), int(warmup)
        self.__dict__["_graphs"] = {}              # plain dict: the wrapped module must not be registered once per shape

    @property
    def aux_loss(self) -> torch.Tensor:
        return self.moe.aux_loss

    @property
    def last_aux(self):
        return self.moe.last_aux

    @property
    def cached_shapes(self):
        return sorted(k_[0] for k_ in self._graphs)

    def forward(self, x: torch.Tensor, context: Optional[torch.Tensor] = None) -> torch.Tensor:
        moe = self.moe
        if not torch.is_grad_enabled() or moe.ep_size > 1 or not x.is_cuda:
            return moe(x, context)
        key = (tuple(x.shape), x.dtype, None if context is None else tuple(context.shape))
        g = self._graphs.get(key)
        if g is None:
            if len(self._graphs) >= self.max_shapes:
                return moe(x, context)
            g = GraphedMoE(moe, x.detach(), context, warmup=self._warmup)
            self._graphs[key] = g
        return g(x, context)
