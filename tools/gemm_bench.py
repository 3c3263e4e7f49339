"""Row-mode GEMM stages alone (CUDA events, L2 flushed):  python tools/gemm_bench.py E d F rows_per_expert
Tells the store-bound epilogues apart: up with and without the pre-activation store, down, dgrad_down, dgrad_up."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "multimodal-moe_b200"))
import torch
import b200moe
from b200moe import ops

E, d, F, n = (int(a) for a in sys.argv[1:5])
b200moe.load_library()
R = E * n
offs = (torch.arange(E + 1, dtype=torch.int32) * n).cuda()
meta = torch.tensor([R, R // 128, 0, R, 0, 0, 0, 0], dtype=torch.int32).cuda()
te = torch.arange(R // 128, dtype=torch.int32).div(n // 128, rounding_mode="floor").int().cuda()
g = torch.Generator(device="cuda").manual_seed(0)
xp = torch.randn(R, d, device="cuda", generator=g).to(torch.bfloat16)
w1 = (torch.randn(E, F, d, device="cuda", generator=g) * d ** -0.5).to(torch.bfloat16)
w2 = (torch.randn(E, d, F, device="cuda", generator=g) * F ** -0.5).to(torch.bfloat16)
b1 = torch.zeros(E, F, device="cuda")
b2 = torch.zeros(E, d, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
pad = 256 if n % 256 == 0 else 128
ACT = int(os.environ.get("GEMM_BENCH_ACT", "0"))           # 0 gelu, 1 relu (cheap epilogue: tells feed-bound from epilogue-bound)
print("env:", {k: v for k, v in os.environ.items() if k.startswith(("B200MOE_GEMM", "GEMM_BENCH"))}, flush=True)


def timed(name, fn, flops, nbytes):
    ts = []
    for i in range(8):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    us = ts[len(ts) // 2]
    print(f"{name:22s} {us:8.1f} us  {flops / us / 1e6:7.0f} TFLOP/s  {nbytes / us / 1e3:6.0f} GB/s (algorithmic)", flush=True)


fl = 2.0 * R * d * F
pre, h = ops.ffn_up(xp, w1, b1, meta, te, save_pre=True, row_pad=pad, activation=ACT)
timed("up (pre + h)", lambda: ops.ffn_up(xp, w1, b1, meta, te, save_pre=True, row_pad=pad, activation=ACT), fl, R * d * 2 + 2 * R * F * 2)
timed("up (h only)", lambda: ops.ffn_up(xp, w1, b1, meta, te, save_pre=False, row_pad=pad, activation=ACT), fl, R * d * 2 + R * F * 2)
timed("down", lambda: ops.ffn_down(h, w2, b2, meta, te, row_pad=pad), fl, R * F * 2 + R * d * 2)
dyp = xp
timed("dgrad_down", lambda: ops.ffn_dgrad_down(dyp, w2, pre, meta, te, row_pad=pad, activation=ACT), fl, R * d * 2 + 2 * R * F * 2)
timed("dgrad_up", lambda: ops.ffn_dgrad_up(h, w1, meta, te, row_pad=pad), fl, R * F * 2 + R * d * 2)
dh = torch.randn(R, F, device="cuda", generator=g).to(torch.bfloat16)
dw2 = torch.empty(E, d, F, device="cuda")
dw1 = torch.empty(E, F, d, device="cuda")
timed("wgrad2 (dw2 + dw1)", lambda: ops.ffn_wgrad2(dyp, h, dh, xp, offs, E, out0=dw2, out1=dw1), 2 * fl,
      2 * (R * d * 2 + R * F * 2) + 2 * E * d * F * 4)
