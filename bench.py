"""bench.py — MoE layer fwd+bwd tokens/s on B200 (BASELINE.json metric), with roofline and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload A|C|D] [--no-sweep]

One "step" = one forward + backward pass of the context-gated MoE layer over one batch of synthetic encoder
tokens.  N=1 runs BASELINE config[1]'s layer shape (E=8, top-2, d=256, ffn=1024, 8x400 tokens + context bins,
bf16); N>1 (under torchrun) runs BASELINE config[3] (E=64, top-2, d=1024, ffn=4096) expert-parallel with 16384
tokens per rank.  Prints ONE JSON line on rank 0 (contract in the task statement): value = tokens/s with inputs
resident in HBM (CUDA events, max over ranks, L2 flushed between steps); e2e = same metric through the public
nn.Module call with pinned HOST inputs and a device->host read of the loss inside the timed region; roofline for the
dominant kernel family (the tcgen05 grouped GEMMs) from a per-call CUDA-event pass; cpu_baseline = the fp32 oracle
on this box's host cores.  `--impl reference` times the oracle itself (the reference has no implementation of this
path) on the same config.
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
}
CONTEXT_VOCAB = (6, 10, 4, 6, 4)   # oracle.moe_ref.CONTEXT_VOCAB_SIZES (add_solar_context_bins.py:90-107 + csv)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return dict(hbm=j["hbm_gbs"], tf_burst=j["bf16_tflops"], tf_sust=j["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


def ncu_traffic(workload):
    """DRAM bytes per launch of the GEMM family from the committed ncu capture (profiles/r01_traffic.json), or None."""
    try:
        j = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))[workload]
        return j["gemm_family_dram_bytes_per_step"] / j["gemm_launches_per_step"]
    except Exception:  # noqa: BLE001
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


def make_inputs(w, seed=1234, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(w["B"], w["HW"], w["d"], generator=g).to(torch.bfloat16)
    go = torch.randn(w["B"], w["HW"], w["d"], generator=g).to(torch.bfloat16)
    ctx = torch.stack([torch.randint(0, v, (w["B"],), generator=g) for v in CONTEXT_VOCAB], 1)
    return x.to(device), go.to(device), ctx.to(device)


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
        "combine": 2 * A * d + 2 * T * d,                           # read kept rows, write y
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


def run_reference(args, w):
    """CPU arm: the fp32 oracle (port; the reference repo has no MoE implementation) on all host threads."""
    from oracle import moe_ref as R
    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(0)
    m = R.ContextMoERef(w["d"], w["F"], w["E"], w["k"], w["cf"], context_vocab_sizes=CONTEXT_VOCAB)
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

    for _ in range(args.warmup):
        step()
    ts = []
    for _ in range(args.steps):
        t0 = time.perf_counter(); step(); ts.append(time.perf_counter() - t0)
    ms = 1e3 * sum(ts) / len(ts)
    return T / (ms * 1e-3), ms, torch.get_num_threads(), f"{sample_B}x{w['HW']} tokens of the workload, fp32, fwd+bwd"


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------
GEMM_CALLS = ("ffn_up", "ffn_down", "ffn_dgrad_down", "ffn_dgrad_up", "ffn_wgrad")


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
                  use_graph=True):
    from b200moe import _lib
    m = build_layer(w, device, ep_group)
    x_h, go_h, ctx_h = make_inputs(w, seed=1234 + (torch.distributed.get_rank() if dist_on else 0))
    x, go, ctx = x_h.to(device), go_h.to(device), ctx_h.to(device)
    T = w["B"] * w["HW"]
    flush = L2Flusher(device)
    xg = x.clone().requires_grad_(True)

    def step():
        m._wcache.clear()                     # weights change every training step: re-cast fp32 -> bf16 each step
        m.zero_grad(set_to_none=True)
        xg.grad = None
        y = m(xg, ctx)
        torch.autograd.backward([y, m.aux_loss], [go, torch.ones_like(m.aux_loss)])

    # (1) eager: every kernel launched from Python through the C ABI (host launch latency included)
    _lib.Counters.reset()
    _lib.load().b200moe_launch_count(1)                      # counted inside the library, at every launch site
    ms_eager = time_steps(step, steps, warmup, flush, dist_on)
    launches = int(_lib.load().b200moe_launch_count(0)) // (steps + warmup)
    res = {"ms_eager": ms_eager, "T": T, "launches_per_step": launches,
           "kept": int(m.last_aux["meta"][3]), "dropped": int(m.last_aux["meta"][2])}

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
        # device-side sleep ahead of each step lets the host queue the whole step first, so durations are GPU time
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
        res["e2e_api"] = "GraphedMoE" if "graphed" in calls else "ContextMoE"
        res["e2e_ms"] = res["e2e_ms_graphed"] if "graphed" in calls else res["e2e_ms_eager"]
        res["h2d"] = x_h.numel() * 2 + ctx_h.numel() * 8
        res["d2h"] = 4
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=[None, "A", "C", "D"])
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

    if args.impl == "reference":
        if rank != 0:
            return
        steps, warmup = min(args.steps, 20), min(args.warmup, 5)
        args.steps, args.warmup = max(steps, 1), warmup
        v, ms, cores, sample = run_reference(args, w)
        line = {"impl": "reference", "metric": "moe_layer_fwd_bwd_tokens_per_s", "value": v, "unit": "tokens/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": {"workload": w["name"]},
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
    T_all = res["T"] * world
    value = T_all / (ms_per_step * 1e-3)

    line = {"metric": "moe_layer_fwd_bwd_tokens_per_s", "value": value, "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": w["name"], "tokens_per_step": T_all, "l2": "flushed between steps (256 MB write)",
                       "parallelism": "single GPU" if world == 1 else f"expert-parallel ep{world}",
                       "kept_assignments": res["kept"], "dropped_assignments": res["dropped"]},
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
    if "gemm_ms" in res and res["gemm_ms"] > 0:
        fl = flops_per_step(w, res["kept"]) if world == 1 else None
        if fl is not None:
            ach = fl / (res["gemm_ms"] * 1e-3) / 1e12
            peak = pk["tf_burst"] if ms_per_step < 1.0 else pk["tf_sust"]
            line["roofline"] = {"bound": "tensor", "kernel": "grouped_gemm_kernel (6 launches: up, down, 2 dgrad, 2 wgrad)",
                                "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                                "peak_source": pk["src"] + (" burst" if ms_per_step < 1.0 else " sustained"),
                                "flops_per_step": fl, "gemm_ms_per_step": res["gemm_ms"],
                                "launches": res.get("gemm_launches"), "traffic": ncu_traffic(wl_name)}
        line["per_call_ms"] = res["per_call_ms"]
        if world == 1:
            line["hbm_kernels"] = hbm_kernels(w, res["T"], res["kept"], res["per_call_ms"], pk["hbm"])

    if rank == 0 and world == 1:
        # CPU baseline: the oracle on this box's host cores, bounded sample
        ra = argparse.Namespace(steps=5, warmup=2)
        v, ms, cores, sample = run_reference(ra, w)
        line["cpu_baseline"] = {"value": v, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample,
                                "ms_per_step": ms}
        if not args.no_sweep:
            sweep = []
            for name, B, HW in (("A", 8, 400), ("A", 164, 400), ("C", 16, 920), ("C", 71, 923), ("D", 16, 1024)):
                ww = dict(WORKLOADS[name], B=B, HW=HW)
                try:
                    r = measure_layer(ww, min(args.steps, 10), 3, device, want_e2e=False)
                    mean_ms = sum(r["ms"]) / len(r["ms"])
                    fl = flops_per_step(ww, r["kept"])
                    pkk = pk["tf_burst"] if mean_ms < 1.0 else pk["tf_sust"]
                    sweep.append({"workload": name, "tokens": B * HW, "ms_per_step": mean_ms, "cuda_graph": r["graph"],
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
        # the same workload on ONE GPU (all experts local, this rank's tokens), measured by rank 0 in the same run, so
        # that value / (n_gpus * scaling_reference.value) is the weak-scaling efficiency on identical per-GPU work
        if rank == 0:
            try:
                r1 = measure_layer(w, min(args.steps, 10), 3, device, want_e2e=False, want_roofline=False)
                ms1 = sum(r1["ms"]) / len(r1["ms"])
                line["scaling_reference"] = {"n_gpus": 1, "workload": w["name"], "ms_per_step": ms1,
                                             "value": r1["T"] / (ms1 * 1e-3), "cuda_graph": r1["graph"]}
            except Exception as ex:  # noqa: BLE001
                line["scaling_reference"] = {"error": repr(ex)}
        dist.barrier()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist_on:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
