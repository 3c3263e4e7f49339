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
