"""bench.py — MoE layer fwd+bwd tokens/s on B200 (BASELINE.json metric), with roofline and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload A|C|D|E] [--no-sweep]

One "step" = one forward + backward pass of the context-gated MoE layer over one batch of synthetic encoder
tokens.  N=1 runs BASELINE configs[1]'s layer shape (E=8, top-2, d=256, ffn=1024, 8x400 tokens + context bins,
bf16); N>1 (under torchrun) runs BASELINE configs[3] (E=64, top-2, d=1024, ffn=4096) expert-parallel with 16384
tokens per rank.  Prints ONE JSON line on rank 0 (contract in the task statement):
  value        tokens/s with inputs resident in HBM (CUDA events, max over ranks, L2 flushed between steps)
  e2e          the same metric through the public nn.Module call with pinned HOST inputs and a device->host read of the
               loss inside the timed region
  roofline     the dominant kernel family (the tcgen05 grouped GEMMs) from a per-call CUDA-event pass, against the
               measured bf16 peak; at N>1 also `nvlink` (bytes that crossed the links, their time, the peer-copy
               bandwidth measured in the same run)
  scaling_same_workload   configs[3]'s shape on ONE GPU measured in the same run, so that weak-scaling efficiency can
               be recomputed from numbers of this line (the driver's N=1 line is configs[1], a different workload)
  ep_parity    (N>1) one step of configs[3]'s shape through the expert-parallel layer against the fp32 oracle on the
               concatenated batch (rank 0, CPU), outside the timed region
  cpu_baseline the fp32 oracle on this box's host cores (N=1).
`--impl reference` times the oracle itself (the reference has no implementation of this path) on the
same config.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "multimodal-moe_b200"))

import torch  # noqa: E402

WORKLOADS = {
    # name: (B, HW, d, F, E, k, capacity_factor)
    "A": dict(B=8, HW=400, d=256, F=1024, E=8, k=2, cf=1.25,
              name="configs[1]: RT-DETR-R18 encoder MoE-FFN layer, 8x400 tokens, d=256 ffn=1024 E=8 top-2 cf=1.25"),
    "C": dict(B=16, HW=920, d=512, F=2048, E=16, k=2, cf=1.25,
              name="configs[2]: E=16 top-2 d=512 ffn=2048 cf=1.25, 16x920 tokens"),
    "D": dict(B=16, HW=1024, d=1024, F=4096, E=64, k=2, cf=1.25,
              name="configs[3]: E=64 top-2 d=1024 ffn=4096 cf=1.25, 16384 tokens per rank, expert-parallel"),
    "E": dict(B=16, HW=920, d=512, F=2048, E=16, k=2, cf=1.25,
              name="configs[4]: context-gated MoE neck, 16x1280x736 frames -> 16x920 tokens per rank, d=512 ffn=2048 "
                   "E=16 top-2 cf=1.25 (YOLO26m itself is not in the reference: synthetic neck tokens, SURVEY.md §8(d))"),
}
CONTEXT_FIELDS = ("solar_context_bin", "scraped_weather", "time_of_day", "road_type", "road_condition")
CONTEXT_VOCAB = (6, 10, 4, 6, 4)   # b200moe.context.CONTEXT_VOCAB_SIZES (add_solar_context_bins.py:90-107 + the csv)
# Empirical counts of outputs/analysis/camera/detection/context_field_frequencies_final.csv:2-26 (context
# ids are drawn from these), as id -> count with id 0 = "missing" (never observed) and the remaining ids in the order of
# b200moe.context.CONTEXT_VOCAB (solar_labels order / sorted category name).
CONTEXT_COUNTS = {
    "solar_context_bin": {"night(<-6)": 19006, "twilight(-6..0)": 3647, "low_sun(0..15)": 16332, "mid_sun(15..45)": 41417,
                          "high_sun(>45)": 19597},
    "scraped_weather": {"partly-cloudy-day": 29153, "cloudy": 18991, "clear-day": 16441, "rain": 15715,
                        "partly-cloudy-night": 11124, "clear-night": 3999, "fog": 2117, "snow": 1972, "wind": 487},
    "time_of_day": {"day": 77346, "night": 19006, "twilight": 3647},
    "road_type": {"city": 50061, "arterial-urban": 23802, "highway": 11061, "arterial-rural": 9925, "smaller-rural": 5150},
    "road_condition": {"normal": 77774, "wet": 15023, "snow": 7202},
}
SOLAR_ORDER = ["night(<-6)", "twilight(-6..0)", "low_sun(0..15)", "mid_sun(15..45)", "high_sun(>45)"]


def context_probs():
    """Per field: probability of every id (0 = missing)."""
    out = []
    for f in CONTEXT_FIELDS:
        names = SOLAR_ORDER if f == "solar_context_bin" else sorted(CONTEXT_COUNTS[f])
        c = torch.tensor([0.0] + [float(CONTEXT_COUNTS[f][n]) for n in names])
        out.append(c / c.sum())
    return out


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return dict(hbm=j["hbm_gbs"], tf_burst=j["bf16_tflops"], tf_sust=j["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


def ncu_traffic(key):
    """DRAM bytes per launch of the GEMM family, REPLAYED from the committed ncu --set full capture of this round
    (profiles/r02_traffic.json, keyed by workload and token count), or None when no capture exists for the key."""
    for name in ("r02_traffic.json", "r01_traffic.json"):
        try:
            j = json.load(open(os.path.join(ROOT, "profiles", name)))[key]
            return {"bytes_per_launch": j["gemm_family_dram_bytes_per_step"] / j["gemm_launches_per_step"],
                    "source": f"replayed from profiles/{name} (ncu --set full), not measured by this run"}
        except Exception:  # noqa: BLE001
            continue
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:  # noqa: BLE001
                pass
        busy = sorted(sm)[len(sm) // 2:] if sm else []      # upper half ~ samples taken under load
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_inputs(w, seed=1234, device="cpu", context="empirical"):
    """unit-normal tokens and output gradient; context ids per image from the empirical frequencies
    of the reference's csv (or uniform over the vocabulary)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(w["B"], w["HW"], w["d"], generator=g).to(torch.bfloat16)
    go = torch.randn(w["B"], w["HW"], w["d"], generator=g).to(torch.bfloat16)
    if context == "uniform":
        ctx = torch.stack([torch.randint(0, v, (w["B"],), generator=g) for v in CONTEXT_VOCAB], 1)
    else:
        ctx = torch.stack([torch.multinomial(p, w["B"], replacement=True, generator=g) for p in context_probs()], 1)
    return x.to(device), go.to(device), ctx.to(device)


def skew_router(m):
    """SURVEY.md §8(d) skewed-router case: +2.0 on the logits of experts 0 .. E/8-1 for every token, through the
    additive context path of A.2 (field 0's embedding rows all equal, its slice of ctx_proj carries the offset), so
    that the capacity limit really drops assignments."""
    cd = m.context_dim
    n_hot = max(1, m.num_experts // 8)
    with torch.no_grad():
        m.ctx_embed[0].weight.fill_(cd ** -0.5)
        m.ctx_proj.weight[:, :cd] = 0.0
        m.ctx_proj.weight[:n_hot, :cd] = 2.0 * cd ** -0.5


def flops_per_step(w, kept_rows):
    return 12.0 * kept_rows * w["d"] * w["F"]      # fwd 4AdF + bwd 8AdF


def hbm_kernels(w, T, kept, per_call_ms, peak_gbs):
    """Algorithmic HBM bytes (A = kept assignments, bf16 = 2 B) of the bandwidth-bound kernels and
    what the per-call CUDA-event durations make of them."""
    d, F, E, k = w["d"], w["F"], w["E"], w["k"]
    A = kept
    algo = {
        "router_fwd": T * d * 2 + T * E * 4 + T * k * 8,            # read x; write logits, top-k indices and gates
        "permute": 4 * A * d,                                       # gather read + write
        "combine": 2 * A * d + 2 * T * d,                           # (separate kernels only; else inside ffn_down's tail)
        "combine_bwd": 6 * A * d,                                   # read dy (gathered) and yo, write dyp
        "unpermute_bwd": 2 * A * d + 2 * T * d,                     # read dxp, write dx
        "colsum": 2 * A * (d + F),                                  # read dyp and dhpre
        "prepare_weights": 6 * 2 * E * d * F,                       # fp32 read + bf16 write of w1, w2
    }
    out = {}
    for name, nbytes in algo.items():
        ms = per_call_ms.get(name)
        if ms:
            gbs = nbytes / (ms * 1e-3) / 1e9
            out[name] = {"bytes": nbytes, "ms": round(ms, 5), "gbs": round(gbs, 1), "frac": round(gbs / peak_gbs, 3)}
    return out


def line_config(w, world, T_all, extra=None):
    """`config` of the JSON line: the same keys in the repo arm and in the reference arm."""
    c = {"workload": w["name"], "tokens_per_step": T_all, "d_model": w["d"], "dim_feedforward": w["F"],
         "num_experts": w["E"], "top_k": w["k"], "capacity_factor": w["cf"],
         "context": "ids drawn from the empirical frequencies of context_field_frequencies_final.csv",
         "parallelism": "single GPU" if world == 1 else f"expert-parallel ep{world}"}
    if extra:
        c.update(extra)
    return c


# ---------------------------------------------------------------------------------------------------------------
# CPU arm (the only place besides tests/ and smoke() that executes oracle/)
# ---------------------------------------------------------------------------------------------------------------
def run_reference(steps, warmup, w, cf="workload"):
    """The fp32 oracle (port; the reference repo has no MoE implementation) on all host threads, fwd+bwd."""
    from oracle import moe_ref as R
    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(0)
    cfv = w["cf"] if cf == "workload" else cf
    m = R.ContextMoERef(w["d"], w["F"], w["E"], w["k"], cfv, context_vocab_sizes=CONTEXT_VOCAB)
    # bounded sample: the full batch for A; a slice of the batch for the big configs (same per-token work)
    sample_B = w["B"] if w["B"] * w["HW"] <= 4096 else max(1, 4096 // w["HW"])
    x, go, ctx = make_inputs(dict(w, B=sample_B))
    x, go = x.float(), go.float()
    T = x.shape[0] * x.shape[1]

    def step():
        m.zero_grad(set_to_none=True)
        xx = x.clone().requires_grad_(True)
        y = m(xx, ctx)
        ((y * go).sum() + m.aux_loss).backward()

    for _ in range(warmup):
        step()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter(); step(); ts.append(time.perf_counter() - t0)
    ms = 1e3 * sum(ts) / len(ts)
    return T / (ms * 1e-3), ms, torch.get_num_threads(), f"{sample_B}x{w['HW']} tokens of the workload, fp32, fwd+bwd"


def run_dense_anchor(steps, warmup, w):
    """Sanity anchor of SURVEY.md §8(d): the dense FFN the layer replaces (hybrid_encoder.py:129-131,157, restated in
    oracle.dense_ffn and pinned to the vendored layer by tests/golden) on the same tokens, fp32, fwd+bwd, all threads."""
    from oracle import moe_ref as R
    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(0)
    d, F = w["d"], w["F"]
    l1, l2 = torch.nn.Linear(d, F), torch.nn.Linear(F, d)
    x, go, _ = make_inputs(w)
    x, go = x.float().reshape(-1, d), go.float().reshape(-1, d)

    def step():
        for p in list(l1.parameters()) + list(l2.parameters()):
            p.grad = None
        xx = x.clone().requires_grad_(True)
        (R.dense_ffn(xx, l1.weight, l1.bias, l2.weight, l2.bias) * go).sum().backward()

    for _ in range(warmup):
        step()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter(); step(); ts.append(time.perf_counter() - t0)
    ms = 1e3 * sum(ts) / len(ts)
    return {"value": x.shape[0] / (ms * 1e-3), "unit": "tokens/s", "ms_per_step": ms,
            "what": "dense FFN of hybrid_encoder.py:157 (restatement), fp32 CPU, fwd+bwd: anchor, not a baseline"}


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------
GEMM_CALLS = ("ffn_up", "ffn_up_gather", "ffn_down", "ffn_dgrad_down", "ffn_dgrad_up", "ffn_wgrad", "ffn_wgrad2",
              "ffn_wgrad_gather")
NVLINK_CALLS = ("ep_p2p_dispatch", "ep_p2p_combine", "ep_p2p_combine_bwd", "ep_p2p_pull_rows")


def build_layer(w, device, ep_group=None):
    import b200moe
    torch.manual_seed(0)
    m = b200moe.ContextMoE(w["d"], w["F"], w["E"], w["k"], w["cf"], context_vocab_sizes=CONTEXT_VOCAB,
                           ep_group=ep_group)
    return m.to(device)


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 between timed steps (outside the timed region)."""

    def __init__(self, device, mb=256):
        self.buf = torch.empty(mb * 1024 * 1024 // 4, dtype=torch.float32, device=device)

    def __call__(self):
        self.buf.fill_(1.0)


def time_steps(step_fn, steps, warmup, flush, dist_on):
    """K individually event-bracketed steps (L2 flushed in between); returns per-step ms list (this rank)."""
    import torch.distributed as dist
    for _ in range(warmup):
        flush(); step_fn()
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    evs = []
    for _ in range(steps):
        flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); step_fn(); b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    if dist_on:
        dist.barrier()
    return [a.elapsed_time(b) for a, b in evs]


def measure_layer(w, steps, warmup, device, ep_group=None, dist_on=False, want_e2e=True, want_roofline=True,
                  use_graph=True, skew=False, recast_weights=True):
    from b200moe import _lib
    m = build_layer(w, device, ep_group)
    if skew:
        skew_router(m)
    x_h, go_h, ctx_h = make_inputs(w, seed=1234 + (torch.distributed.get_rank() if dist_on else 0))
    x, go, ctx = x_h.to(device), go_h.to(device), ctx_h.to(device)
    T = w["B"] * w["HW"]
    flush = L2Flusher(device)
    xg = x.clone().requires_grad_(True)
    one = torch.ones((), dtype=torch.float32, device=device)     # upstream gradient of the auxiliary loss

    def step():
        if recast_weights:
            m._wcache.clear()                 # weights change every training step: re-cast fp32 -> bf16 each step
        m.zero_grad(set_to_none=True)
        xg.grad = None
        y = m(xg, ctx)
        torch.autograd.backward([y, m.aux_loss], [go, one])

    # (1) eager: every kernel launched from Python through the C ABI (host launch latency included)
    _lib.Counters.reset()
    _lib.load().b200moe_launch_count(1)                      # counted inside the library, at every launch site
    ms_eager = time_steps(step, steps, warmup, flush, dist_on)
    launches = int(_lib.load().b200moe_launch_count(0)) // (steps + warmup)
    res = {"ms_eager": ms_eager, "T": T, "launches_per_step": launches,
           "kept": int(m.last_aux["meta"][3]), "dropped": int(m.last_aux["meta"][2])}
    if dist_on:
        # rows of this rank that live on another rank's experts: the NVLink payload of dispatch / combine
        EL = w["E"] // torch.distributed.get_world_size()
        idx, roa = m.last_aux["topk_idx"], m.last_aux["row_of_assignment"]
        remote = ((idx // EL) != torch.distributed.get_rank()) & (roa >= 0)
        res["remote_rows"] = int(remote.sum())
        res["transport"] = "p2p" if "_p2p_arena" in m.__dict__ else "nccl"

    # (2) the same step captured once into a CUDA graph and replayed (no host sync anywhere in the layer)
    res["ms"] = ms_eager
    res["graph"] = False
    if not dist_on and use_graph:
        try:
            graph = torch.cuda.CUDAGraph()
            m.zero_grad(set_to_none=True)
            xg.grad = None
            m.aux_loss, m.last_aux = None, {}        # drop the eager autograd graph (its AccumulateGrad nodes
            side = torch.cuda.Stream()               # belong to the default stream) and warm up on a side stream
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    step()
                m.zero_grad(set_to_none=True)
                xg.grad = None
                m.aux_loss, m.last_aux = None, {}
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            with torch.cuda.graph(graph):
                step()
            torch.cuda.synchronize()
            res["ms"] = time_steps(graph.replay, steps, warmup, flush, dist_on)
            res["graph"] = True
            del graph
        except Exception as ex:  # noqa: BLE001
            res["graph_error"] = repr(ex)

    if want_roofline:
        # (3) instrumented eager pass: every C-ABI call bracketed by CUDA events on the launching stream; a ~3 ms
        # device-side sleep ahead of each step lets the host queue the whole step first, so durations are GPU time.
        # The weight-preparation side stream is off while recording (the layer sees _lib.PROFILING), so the router /
        # plan / permute numbers are those of kernels running alone.
        torch.cuda.synchronize()
        _lib.Counters.reset(timing=True)
        _lib.layer_profile(True)              # events around every kernel inside b200moe_layer_fwd / _bwd
        for _ in range(steps):
            torch.cuda._sleep(6_000_000)
            flush(); step()
        torch.cuda.synchronize()
        dur = {k_.replace("b200moe_", ""): v for k_, v in _lib.Counters.durations_ms().items()
               if k_ not in ("b200moe_layer_fwd", "b200moe_layer_bwd")}
        for k_, v in _lib.layer_profile_durations().items():
            dur.setdefault(k_, []).extend(v)
        _lib.layer_profile(False)
        _lib.Counters.reset()
        per_step = {k_: sum(v) / steps for k_, v in dur.items()}
        res["per_call_ms"] = {k_: round(v, 5) for k_, v in sorted(per_step.items())}
        res["gemm_ms"] = sum(per_step.get(n, 0.0) for n in GEMM_CALLS)
        res["gemm_launches"] = sum(len(dur.get(n, [])) for n in GEMM_CALLS) // steps
        res["nvlink_ms"] = {n: per_step[n] for n in NVLINK_CALLS if n in per_step}

    if want_e2e:
        xp_h, ctxp_h = x_h.pin_memory(), ctx_h.pin_memory()
        m.zero_grad(set_to_none=True)
        m.aux_loss, m.last_aux = None, {}
        # the public call for fixed-shape training: b200moe.GraphedMoE (forward and backward replayed from CUDA graphs
        # behind the module's autograd interface); the expert-parallel pass is launched eagerly
        calls = {"eager": m}
        if not dist_on and use_graph:
            import b200moe
            calls["graphed"] = b200moe.GraphedMoE(m, x, ctx)

        # input pipeline of a training loop: pinned host batch -> device on a copy stream, one step ahead (every timed
        # step contains one host->device copy of a full batch, the layer's forward + backward, and the loss read back)
        copy_stream = torch.cuda.Stream(device=device)
        bufs = [(torch.empty_like(x), torch.empty_like(ctx)) for _ in range(2)]
        evs = [torch.cuda.Event() for _ in range(2)]

        def prefetch(i):
            with torch.cuda.stream(copy_stream):
                bufs[i % 2][0].copy_(xp_h, non_blocking=True)
                bufs[i % 2][1].copy_(ctxp_h, non_blocking=True)
                evs[i % 2].record(copy_stream)

        def e2e_step(call, i, pipelined):
            m.zero_grad(set_to_none=True)
            if pipelined:
                torch.cuda.current_stream().wait_event(evs[i % 2])
                xd = bufs[i % 2][0].detach().requires_grad_(True)
                cd = bufs[i % 2][1]
            else:                                           # same-stream copies just before the step
                xd = xp_h.to(device, non_blocking=True).requires_grad_(True)
                cd = ctxp_h.to(device, non_blocking=True)
            y = call(xd, cd)
            loss = (y.float() * go).sum() + call.aux_loss
            loss.backward()
            if pipelined:
                prefetch(i + 1)                             # the other buffer: its last reader finished a step ago
            return float(loss.item())                       # device -> host read of the step's result

        if not dist_on and use_graph:
            # the whole step (forward, loss, backward) as one CUDA graph: b200moe.GraphedStep.  Pinned host inputs are
            # copied straight into the captured buffers; the loss is read back every step.
            import b200moe
            gstep = b200moe.GraphedStep(m, lambda y, g_: (y.float() * g_).sum(), x, ctx, go)
            torch.cuda.synchronize()
            for i in range(warmup):
                float(gstep(xp_h, ctxp_h).item())
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for i in range(steps):
                float(gstep(xp_h, ctxp_h).item())
            b.record()
            torch.cuda.synchronize()
            res["e2e_ms_step_inline"] = a.elapsed_time(b) / steps
            del gstep
            # the same with the copy of batch i+1 on a copy stream under step i (b200moe.PipelinedGraphedStep): every
            # timed step still contains one host->device copy of a full batch and the loss read-back
            pstep = b200moe.PipelinedGraphedStep(m, lambda y, g_: (y.float() * g_).sum(), x, ctx, go)
            pstep.prefetch(xp_h, ctxp_h)
            for i in range(warmup):
                l_ = pstep(); pstep.prefetch(xp_h, ctxp_h); float(l_.item())
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for i in range(steps):
                l_ = pstep(); pstep.prefetch(xp_h, ctxp_h); float(l_.item())
            b.record()
            torch.cuda.synchronize()
            res["e2e_ms_step_prefetch"] = a.elapsed_time(b) / steps
            del pstep
        for name, call in calls.items():
            for pipelined in (False, True):
                torch.cuda.synchronize()
                prefetch(0)
                for i in range(warmup):
                    e2e_step(call, i, pipelined)
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for i in range(warmup, warmup + steps):
                    e2e_step(call, i, pipelined)
                b.record()
                torch.cuda.synchronize()
                res["e2e_ms_" + name + ("_prefetch" if pipelined else "_inline")] = a.elapsed_time(b) / steps
            res["e2e_ms_" + name] = min(res["e2e_ms_" + name + "_prefetch"], res["e2e_ms_" + name + "_inline"])
        if not dist_on and use_graph:
            # multi-scale training (dataloader.yml:22): the token count changes per step, so one captured shape cannot
            # serve; b200moe.MultiShapeGraphedMoE keeps a graph per shape.  Same rules: pinned host batch in, loss out.
            import b200moe
            ms_ = {}
            scales = [(w["B"], hw) for hw in (225, 400, 625)]
            batches = []
            for B_, hw in scales:
                xb, gb, cb = make_inputs(dict(w, B=B_, HW=hw), seed=77 + hw)
                batches.append((xb.pin_memory(), cb.pin_memory(), gb.to(device)))
            for name, call in (("eager", m), ("multishape_graphed", b200moe.MultiShapeGraphedMoE(m))):
                def ms_step(i, call=call):
                    xh, ch, g_ = batches[i % len(batches)]
                    m.zero_grad(set_to_none=True)
                    xd = xh.to(device, non_blocking=True).requires_grad_(True)
                    y = call(xd, ch.to(device, non_blocking=True))
                    loss = (y.float() * g_).sum() + call.aux_loss
                    loss.backward()
                    return float(loss.item())
                for i in range(max(warmup, 2 * len(batches))):
                    ms_step(i)
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                n_ms = len(batches) * max(steps // len(batches), 1)
                a.record()
                for i in range(n_ms):
                    ms_step(i)
                b.record()
                torch.cuda.synchronize()
                ms_[name] = a.elapsed_time(b) / n_ms
            tok = sum(B_ * hw for B_, hw in scales) / len(scales)
            res["e2e_multiscale"] = {"tokens_per_step": [B_ * hw for B_, hw in scales],
                                     "ms_per_step": {k_: round(v, 5) for k_, v in ms_.items()},
                                     "tokens_per_s": {k_: tok / (v * 1e-3) for k_, v in ms_.items()},
                                     "api": "MultiShapeGraphedMoE (one CUDA graph pair per shape) vs the eager module; "
                                            "shapes cycle 225 / 400 / 625 tokens per image"}
        res["e2e_api"] = "GraphedMoE" if "graphed" in calls else "ContextMoE"
        res["e2e_ms"] = res["e2e_ms_graphed"] if "graphed" in calls else res["e2e_ms_eager"]
        if "e2e_ms_step_inline" in res and res["e2e_ms_step_inline"] < res["e2e_ms"]:
            res["e2e_api"], res["e2e_ms"] = "GraphedStep", res["e2e_ms_step_inline"]
        if "e2e_ms_step_prefetch" in res and res["e2e_ms_step_prefetch"] < res["e2e_ms"]:
            res["e2e_api"], res["e2e_ms"] = "PipelinedGraphedStep", res["e2e_ms_step_prefetch"]
        res["h2d"] = x_h.numel() * 2 + ctx_h.numel() * 8
        res["d2h"] = 4
    return res


def measure_in_situ(device, steps, warmup):
    """BASELINE configs[1] in situ: the S5 feature map [8,256,20,20] goes through the
    flatten / permute of hybrid_encoder.py:299-301 into one post-norm AIFI layer (attention + FFN, :143-161, restated
    here because the vendored package is not on the GPU box) whose dense FFN was upcycled to the MoE
    (b200moe.MoEEncoderLayer.from_dense), under bf16 autocast as the training step runs it (det_engine.py:45).
    Timed with CUDA events: the MoE call alone (forward, brackets around the :157 replacement inside the layer) and
    the whole layer fwd+bwd, dense vs MoE."""
    import copy
    import torch.nn as nn
    import b200moe

    class Layer(nn.Module):                       # hybrid_encoder.py:116-161, post-norm, dropout 0
        def __init__(self, d, nhead, dff):
            super().__init__()
            self.normalize_before = False
            self.self_attn = nn.MultiheadAttention(d, nhead, 0.0, batch_first=True)
            self.linear1, self.linear2 = nn.Linear(d, dff), nn.Linear(dff, d)
            self.norm1, self.norm2 = nn.LayerNorm(d), nn.LayerNorm(d)
            self.dropout, self.dropout1, self.dropout2 = nn.Dropout(0.0), nn.Dropout(0.0), nn.Dropout(0.0)
            self.activation = nn.GELU()

        def forward(self, src, src_mask=None, pos_embed=None):
            q = k = src if pos_embed is None else src + pos_embed
            a, _ = self.self_attn(q, k, value=src, attn_mask=src_mask)
            src = self.norm1(src + self.dropout1(a))
            f = self.linear2(self.dropout(self.activation(self.linear1(src))))
            return self.norm2(src + self.dropout2(f))

    torch.manual_seed(0)
    w = WORKLOADS["A"]
    B, d, Hs = w["B"], w["d"], 20
    dense = Layer(d, 8, w["F"]).to(device)
    moe = b200moe.MoEEncoderLayer.from_dense(copy.deepcopy(dense), num_experts=w["E"], top_k=w["k"],
                                             capacity_factor=w["cf"], context_vocab_sizes=CONTEXT_VOCAB)
    _, _, ctx = make_inputs(w, device=device)
    moe.set_context(ctx)
    feat = torch.randn(B, d, Hs, Hs, device=device)
    pos = torch.randn(1, Hs * Hs, d, device=device)
    go = torch.randn(B, Hs * Hs, d, device=device)
    flush = L2Flusher(device)
    fwd_ev = []
    orig = moe.moe.forward_post_norm

    def bracketed(*a, **k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = orig(*a, **k); e1.record()
        fwd_ev.append((e0, e1))
        return out
    moe.moe.forward_post_norm = bracketed

    def make_step(layer, is_moe):
        def step():
            for p in layer.parameters():
                p.grad = None
            f = feat.clone().requires_grad_(True)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                src = f.flatten(2).permute(0, 2, 1)                    # hybrid_encoder.py:299-301
                out = layer(src, pos_embed=pos)
                loss = (out.float() * go).sum()
                if is_moe:
                    loss = loss + layer.aux_loss
            loss.backward()
        return step

    out = {"shape": f"features [{B},{d},{Hs},{Hs}] -> tokens [{B},{Hs * Hs},{d}], bf16 autocast, eager launches"}
    for name, layer, is_moe in (("dense", dense, False), ("moe", moe, True)):
        fwd_ev.clear()
        ms = time_steps(make_step(layer, is_moe), steps, warmup, flush, False)
        out[f"layer_fwd_bwd_ms_{name}"] = sum(ms) / len(ms)
    calls = fwd_ev[-steps:]
    out["moe_call_fwd_ms"] = sum(a.elapsed_time(b) for a, b in calls) / max(len(calls), 1)
    out["moe_call"] = "LayerNorm(residual + MoE(x)) fused path (forward_post_norm), host launch gaps included"
    out["tokens_per_s_layer_moe"] = B * Hs * Hs / (out["layer_fwd_bwd_ms_moe"] * 1e-3)
    return out


def measure_nvlink_peak(device, group):
    """Per-direction peer-copy bandwidth of this box, measured in this run: every rank copies 256 MiB into the
    symmetric buffer of rank+1 at the same time (all links busy in both directions, as during dispatch / combine)."""
    import torch.distributed as dist
    try:
        import torch.distributed._symmetric_memory as symm
        W, rank = dist.get_world_size(group), dist.get_rank(group)
        n = 128 * 1024 * 1024
        mem = symm.empty(n, dtype=torch.bfloat16, device=device)
        hdl = symm.rendezvous(mem, group)
        peer = hdl.get_buffer((rank + 1) % W, (n,), torch.bfloat16)
        src = torch.ones(n, dtype=torch.bfloat16, device=device)
        best = 0.0
        for i in range(6):
            hdl.barrier(0)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); peer.copy_(src); b.record()
            torch.cuda.synchronize()
            if i >= 2:
                best = max(best, 2.0 * n / (a.elapsed_time(b) * 1e-3) / 1e9)
        hdl.barrier(0)
        t = torch.tensor([best], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
        return {"peak_gbs_measured": float(t[0]), "how": "torch copy_ of 256 MiB into the next rank's symmetric buffer, "
                "all ranks at once, best of 4, min over ranks"}
    except Exception as ex:  # noqa: BLE001
        return {"peak_gbs_measured": None, "how": f"unavailable: {ex!r}"}


def ep_parity_check(w, device, group, tokens_per_rank=1024):
    """Outside the timed region: one fwd+bwd of the workload's layer shape through the expert-parallel layer, against the
    fp32 oracle (rank 0, CPU) on the concatenated rank-major batch and on the kernel's own fp32 logits (A.11), plus the
    oracle's per-rank `global_offset` formulation of the capacity rule (A.5)."""
    import torch.distributed as dist
    import b200moe
    W, rank = dist.get_world_size(group), dist.get_rank(group)
    d, F, E, k, cf = w["d"], w["F"], w["E"], w["k"], w["cf"]
    B, HW = 2, tokens_per_rank // 2
    T = B * HW
    torch.manual_seed(0)
    m = b200moe.ContextMoE(d, F, E, k, cf, context_vocab_sizes=CONTEXT_VOCAB, ep_group=group)
    with torch.no_grad():                            # bf16-representable expert weights: isolates kernel error
        for p in (m.w1, m.w2):
            p.copy_(p.to(torch.bfloat16).float())
        m.gate.weight[: max(1, E // 8)] *= 3.0       # skew the router so that the global capacity really drops
    m = m.to(device)
    x_h, go_h, ctx_h = make_inputs(dict(w, B=B, HW=HW), seed=77 + rank)
    x = x_h.to(device).requires_grad_(True)
    y = m(x, ctx_h.to(device))
    y.backward(go_h.to(device))       # the aux statistics are per-rank local (A.10): compared on y's gradient alone
    torch.cuda.synchronize()
    transport = "p2p" if "_p2p_arena" in m.__dict__ else "nccl"

    def gather(t):
        t = t.contiguous()
        out = [torch.empty_like(t) for _ in range(W)] if rank == 0 else None
        dist.gather(t, out, dst=0, group=group)
        return torch.cat([o.cpu() for o in out]) if rank == 0 else None

    a = m.last_aux
    g_x, g_go, g_ctx = gather(x.detach()), gather(go_h.to(device)), gather(ctx_h.to(device))
    g_logits, g_idx = gather(a["logits"]), gather(a["topk_idx"])
    g_keep = gather((a["row_of_assignment"] >= 0).to(torch.uint8))
    g_y, g_dx = gather(y.detach()), gather(x.grad)
    dwg = m.gate.weight.grad.clone()
    dist.all_reduce(dwg, group=group)                # replicated parameter: per-rank local sums add up
    dw1_local = m.w1.grad.cpu()
    del m
    out = None
    if rank == 0:
        from oracle import moe_ref as R
        torch.set_num_threads(os.cpu_count() or 1)
        torch.manual_seed(0)
        full = b200moe.ContextMoE(d, F, E, k, cf, context_vocab_sizes=CONTEXT_VOCAB)      # same seed => same global model
        ref = R.ContextMoERef(d, F, E, k, cf, context_vocab_sizes=CONTEXT_VOCAB)
        ref.load_state_dict(full.state_dict())
        del full
        with torch.no_grad():
            for p in (ref.w1, ref.w2):
                p.copy_(p.to(torch.bfloat16).float())
            ref.gate.weight[: max(1, E // 8)] *= 3.0
        xr = g_x.float().reshape(W * B, HW, d).requires_grad_(True)
        yr = ref(xr, g_ctx, logits_value=g_logits)
        (yr * g_go.float().reshape(W * B, HW, d)).sum().backward()

        def rel(u, v):
            return float((u.float() - v.float()).norm() / (v.float().norm() + 1e-12))
        idx_ref = ref.last_aux["topk_idx"]
        keep_ref = ref.last_aux["keep"]
        # per-rank formulation of A.5 (oracle/moe_ref.py assign_slots with global_offset): counts of the lower ranks at
        # the same slot level plus everything at the lower slot levels
        cap = ref.last_aux["capacity"]
        oh = torch.nn.functional.one_hot(idx_ref, E).view(W, T, k, E).sum(1)          # [W,k,E]
        lower_levels = torch.cumsum(oh.sum(0), 0) - oh.sum(0)                          # [k,E]
        ok_offset = True
        for r in range(W):
            goff = lower_levels + oh[:r].sum(0)
            _, keep_r = R.assign_slots(idx_ref[r * T:(r + 1) * T], E, cap, goff)
            ok_offset = ok_offset and torch.equal(keep_r, keep_ref[r * T:(r + 1) * T])
        EL = E // W
        out = {"tokens_per_rank": T, "ranks": W, "transport": transport, "capacity": cap,
               "dropped": int((~keep_ref).sum()),
               "topk_idx_equal": bool(torch.equal(g_idx.long(), idx_ref)),
               "drops_equal": bool(torch.equal(g_keep.bool(), keep_ref)),
               "oracle_global_offset_form_equal": bool(ok_offset),
               "rel_y": rel(g_y.reshape(-1, d), yr.detach().reshape(-1, d)),
               "rel_dx": rel(g_dx.reshape(-1, d), xr.grad.reshape(-1, d)),
               "rel_dw1": rel(dw1_local, ref.w1.grad[:EL]),
               "rel_dwg": rel(dwg.cpu(), ref.gate.weight.grad),
               "tolerance": "routing bit-exact; rel-L2 <= 1e-2 (A.11); gate gradient excludes the per-rank aux loss"}
        out["ok"] = bool(out["topk_idx_equal"] and out["drops_equal"] and ok_offset and
                         max(out["rel_y"], out["rel_dx"], out["rel_dw1"], out["rel_dwg"]) <= 1e-2)
    dist.barrier(group=group)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=[None, "A", "C", "D", "E"])
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-pdl", action="store_true", help="plain stream-ordered launches inside the whole-layer calls")
    ap.add_argument("--fork", type=int, default=1, choices=[0, 1],
                    help="backward-pass side streams (weight / router gradients next to the dgrad chain): 0 off, 1 on")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    wl_name = args.workload or ("A" if world == 1 else "D")
    w = WORKLOADS[wl_name]
    pk = peaks()
    T_all = w["B"] * w["HW"] * world

    if args.impl == "reference":
        if rank != 0:
            return
        # bounded: the CPU arm of a multi-GPU line is context, not a same-size comparison (3 steps there)
        steps = min(args.steps, 3 if world > 1 else 20)
        warmup = min(args.warmup, 1 if world > 1 else 5)
        v, ms, cores, sample = run_reference(max(steps, 1), warmup, w)
        line = {"impl": "reference", "metric": "moe_layer_fwd_bwd_tokens_per_s", "value": v, "unit": "tokens/s",
                "n_gpus": args.gpus, "steps": max(steps, 1), "warmup": warmup, "ms_per_step": ms,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": line_config(w, world, T_all),
                "cpu_baseline": {"value": v, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample},
                "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a B200; there is no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist_on = world > 1
    ep_group = None
    if dist_on:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ.pop("NCCL_DEBUG")          # its banner goes to stdout, which carries the ONE JSON line
        dist.init_process_group("nccl", device_id=device)
        ep_group = dist.group.WORLD

    if args.no_pdl:
        from b200moe import _lib as _l
        _l.load().b200moe_layer_pdl(0)
    if args.fork != 1:
        from b200moe import _lib as _l
        _l.load().b200moe_layer_fork(args.fork)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    res = measure_layer(w, args.steps, args.warmup, device, ep_group, dist_on)
    clocks = sampler.stop() if sampler else None

    # max over ranks of the per-step device times
    ms_t = torch.tensor(res["ms"], device=device)
    e2e_t = torch.tensor([res.get("e2e_ms", 0.0)], device=device)
    if dist_on:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    total_ms = float(ms_t.sum())
    ms_per_step = total_ms / args.steps
    value = T_all / (ms_per_step * 1e-3)

    line = {"metric": "moe_layer_fwd_bwd_tokens_per_s", "value": value, "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": line_config(w, world, T_all, {"l2": "flushed between steps (256 MB write)",
                                                    "kept_assignments": res["kept"],
                                                    "dropped_assignments": res["dropped"]}),
            "gpu_launches": res["launches_per_step"] * args.steps,
            "gpu_launches_per_step": res["launches_per_step"], "clocks": clocks,
            "cuda_graph": res["graph"],
            "eager_ms_per_step": sum(res["ms_eager"]) / len(res["ms_eager"])}
    if "graph_error" in res:
        line["graph_error"] = res["graph_error"]
    if "e2e_ms" in res:
        line["e2e"] = {"value": T_all / (float(e2e_t[0]) * 1e-3), "unit": "tokens/s",
                       "h2d_bytes_per_step": res["h2d"], "d2h_bytes_per_step": res["d2h"], "api": res["e2e_api"],
                       "ms_per_step": {k_[7:]: round(v, 5) for k_, v in res.items() if k_.startswith("e2e_ms_") and k_.count("_") > 2},
                       "pipeline": "best of: inline = inputs copied on the compute stream just before the step; prefetch = "
                                   "batch i+1 copied on a copy stream during step i.  Loss read back every step in both.",
                       "eager_api_value": T_all / (res["e2e_ms_eager"] * 1e-3)}
        if "e2e_multiscale" in res:
            line["e2e"]["multiscale"] = res["e2e_multiscale"]
    if "gemm_ms" in res and res["gemm_ms"] > 0:
        # per rank: kept rows of this rank's tokens at N=1; at N>1 the rows this rank's experts received are, summed
        # over the ranks, the kept rows of all tokens -- the per-rank average is what the FLOP count uses
        kept = res["kept"]
        if dist_on:
            kt = torch.tensor([float(kept)], device=device)
            dist.all_reduce(kt)
            kept = float(kt[0]) / world
            gm = torch.tensor([res["gemm_ms"]], device=device)
            dist.all_reduce(gm, op=dist.ReduceOp.MAX)
            res["gemm_ms"] = float(gm[0])
        fl = flops_per_step(w, kept)
        ach = fl / (res["gemm_ms"] * 1e-3) / 1e12
        peak = pk["tf_burst"] if ms_per_step < 1.0 else pk["tf_sust"]
        key = f"{wl_name}{res['T']}" + (f"_ep{world}" if dist_on else "")
        tr = ncu_traffic(key)
        line["roofline"] = {"bound": "tensor", "kernel": "grouped_gemm_kernel (up, down, dgrad_down, dgrad_up, both weight gradients in one launch)",
                            "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                            "peak_source": pk["src"] + (" burst" if ms_per_step < 1.0 else " sustained"),
                            "flops_per_step": fl, "gemm_ms_per_step": res["gemm_ms"],
                            "per_rank": dist_on, "launches": res.get("gemm_launches"),
                            "traffic": tr["bytes_per_launch"] if tr else None,
                            "traffic_source": tr["source"] if tr else None}
        line["per_call_ms"] = res["per_call_ms"]
        if world == 1:
            line["hbm_kernels"] = hbm_kernels(w, res["T"], res["kept"], res["per_call_ms"], pk["hbm"])

    if dist_on:
        # NVLink: rows of a rank that live on another rank's experts cross the links once per transfer kernel
        nv = measure_nvlink_peak(device, ep_group)
        rr = torch.tensor([float(res.get("remote_rows", 0))], device=device)
        dist.all_reduce(rr, op=dist.ReduceOp.MAX)
        bytes_dir = float(rr[0]) * w["d"] * 2
        nvms = res.get("nvlink_ms", {})
        if nvms:
            t = torch.tensor([nvms.get(n, 0.0) for n in NVLINK_CALLS], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            per = {}
            for n, ms in zip(NVLINK_CALLS, t.tolist()):
                if ms > 0:
                    # combine_bwd stores dyp rows out and loads yo rows in: both directions carry bytes_dir
                    gbs = bytes_dir / (ms * 1e-3) / 1e9
                    per[n] = {"ms": round(ms, 5), "gbs_per_direction": round(gbs, 1),
                              "frac": round(gbs / nv["peak_gbs_measured"], 3) if nv["peak_gbs_measured"] else None}
            tot_ms = sum(v["ms"] for v in per.values())
            nvl = {"bytes_per_dir": bytes_dir, "transfers": per, "ms": tot_ms,
                   "gbs": bytes_dir * len(per) / (tot_ms * 1e-3) / 1e9 if tot_ms else None,
                   "peak_gbs_measured": nv["peak_gbs_measured"], "peak_how": nv["how"], "peak_gbs_nominal": 900.0,
                   "exposed": "dispatch, combine, combine_bwd run on the compute stream (exposed); pull_rows runs on a "
                              "side stream under the weight gradients"}
            nvl["frac"] = nvl["gbs"] / nv["peak_gbs_measured"] if nvl["gbs"] and nv["peak_gbs_measured"] else None
            line.setdefault("roofline", {})["nvlink"] = nvl
        line["ep_transport"] = res.get("transport")
        try:
            line["ep_parity"] = ep_parity_check(w, device, ep_group)
        except Exception as ex:  # noqa: BLE001
            line["ep_parity"] = {"ok": False, "error": repr(ex)}

    if dist_on and w["E"] != WORKLOADS["E"]["E"] and WORKLOADS["E"]["E"] % world == 0:
        # BASELINE.json configs[4] beside the metric's own config: data-parallel tokens (16x920 per rank) + expert-parallel
        # experts (E=16, d=512, ffn=2048), every rank takes part; same timing rules, fewer steps
        we = WORKLOADS["E"]
        try:
            re_ = measure_layer(we, min(args.steps, 10), 3, device, ep_group, dist_on, want_e2e=False, want_roofline=False)
            te = torch.tensor(re_["ms"], device=device)
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
            ms_e = float(te.sum()) / len(re_["ms"])
            line["config_E"] = {"workload": we["name"], "n_gpus": world, "tokens_per_step": we["B"] * we["HW"] * world,
                                "ms_per_step": ms_e, "value": we["B"] * we["HW"] * world / (ms_e * 1e-3),
                                "unit": "tokens/s", "transport": re_.get("transport"), "cuda_graph": re_["graph"],
                                "kept_assignments_rank0": re_["kept"], "dropped_assignments_rank0": re_["dropped"]}
        except Exception as ex:  # noqa: BLE001
            line["config_E"] = {"error": repr(ex)}
        torch.cuda.empty_cache()

    # the same workload (configs[3]) on ONE GPU, measured in the same run: all experts local, one rank's tokens.  The
    # driver's N=1 line is configs[1], so efficiency has to be recomputed from these numbers.
    if rank == 0 and (dist_on or not args.no_sweep):
        wd = WORKLOADS["D"] if not dist_on else w
        try:
            r1 = measure_layer(wd, min(args.steps, 10), 3, device, want_e2e=False, want_roofline=dist_on)
            ms1 = sum(r1["ms"]) / len(r1["ms"])
            sc = {"workload": wd["name"], "n_gpus": 1, "ms_per_step_1gpu": ms1, "value_1gpu": r1["T"] / (ms1 * 1e-3),
                  "cuda_graph": r1["graph"]}
            if dist_on:
                sc["efficiency_same_workload"] = value / (world * sc["value_1gpu"])
                # one GPU prepares (fp32 -> bf16) and writes weight gradients for all E experts, a rank of the
                # expert-parallel run for E/N of them: charge the single GPU only its share of the preparation
                prep = r1.get("per_call_ms", {}).get("prepare_weights", 0.0)
                ms_fair = ms1 - prep * (1.0 - 1.0 / world)
                sc["ms_per_step_1gpu_local_prep"] = ms_fair
                sc["efficiency_vs_local_expert_prep"] = (w["B"] * w["HW"] / (ms_per_step * 1e-3)) / (r1["T"] / (ms_fair * 1e-3))
                sc["note"] = ("value / (n_gpus * value_1gpu); above 1.0 only because a rank casts and updates E/N experts "
                              "(prepare_weights, fp32 dW traffic); efficiency_vs_local_expert_prep removes the cast share")
            line["scaling_same_workload"] = sc
        except Exception as ex:  # noqa: BLE001
            line["scaling_same_workload"] = {"error": repr(ex)}

    if rank == 0 and world == 1:
        # CPU baseline: the oracle on this box's host cores, bounded sample; capacity factor as configured and "no limit"
        v, ms, cores, sample = run_reference(5, 2, w)
        line["cpu_baseline"] = {"value": v, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample,
                                "ms_per_step": ms}
        v2, ms2, _, _ = run_reference(3, 1, w, cf=None)
        line["cpu_baseline"]["capacity_factor_none"] = {"value": v2, "ms_per_step": ms2}
        line["cpu_baseline"]["dense_ffn_anchor"] = run_dense_anchor(5, 2, w)
        try:
            line["in_situ"] = measure_in_situ(device, min(args.steps, 20), 5)
        except Exception as ex:  # noqa: BLE001
            line["in_situ"] = {"error": repr(ex)}
        if not args.no_sweep:
            sweep = []
            # (workload, B, HW, skewed router): T in {3200, 14720, 65536, 262144} per SURVEY.md §8(d) where it fits the
            # shape, the skewed-router case with real drops, and configs[3] on one GPU
            cases = (("A", 8, 400, False), ("A", 8, 400, True), ("A", 164, 400, False), ("A", 655, 400, False),
                     ("C", 16, 920, False), ("C", 16, 920, True), ("C", 71, 923, False), ("C", 284, 923, False),
                     ("D", 16, 1024, False))
            for name, B, HW, skew in cases:
                ww = dict(WORKLOADS[name], B=B, HW=HW)
                try:
                    r = measure_layer(ww, min(args.steps, 10), 3, device, want_e2e=False, skew=skew)
                    mean_ms = sum(r["ms"]) / len(r["ms"])
                    fl = flops_per_step(ww, r["kept"])
                    pkk = pk["tf_burst"] if mean_ms < 1.0 else pk["tf_sust"]
                    sweep.append({"workload": name, "tokens": B * HW, "skewed_router": skew, "ms_per_step": mean_ms,
                                  "cuda_graph": r["graph"], "kept": r["kept"], "dropped": r["dropped"],
                                  "launches_per_step": r["launches_per_step"],
                                  "tokens_per_s": B * HW / (mean_ms * 1e-3), "gemm_ms": r["gemm_ms"],
                                  "gemm_tflops": fl / (r["gemm_ms"] * 1e-3) / 1e12,
                                  "gemm_frac_of_peak": fl / (r["gemm_ms"] * 1e-3) / 1e12 / pkk,
                                  "per_call_ms": r["per_call_ms"],
                                  "hbm_kernels": hbm_kernels(ww, B * HW, r["kept"], r["per_call_ms"], pk["hbm"])})
                except Exception as ex:  # noqa: BLE001
                    sweep.append({"workload": name, "tokens": B * HW, "error": repr(ex)})
                torch.cuda.empty_cache()
            line["sweep"] = sweep
    if dist_on:
        dist.barrier()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist_on:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
