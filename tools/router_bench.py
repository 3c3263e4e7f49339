"""Router forward alone: python tools/router_bench.py T d E [k] [iters]  -> us per launch (CUDA events, L2 flushed) and GB/s.
Also the kernel to point ncu at:  ncu --set full -k regex:router_fwd -c 3 python tools/router_bench.py 65533 512 16"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "multimodal-moe_b200"))
import torch
import b200moe
from b200moe import ops

T, d, E = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
k = int(sys.argv[4]) if len(sys.argv) > 4 else 2
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 20
b200moe.load_library()
g = torch.Generator().manual_seed(0)
x = torch.randn(T, d, generator=g).to(torch.bfloat16).cuda()
wg = (torch.randn(E, d, generator=g) * d ** -0.5).cuda()
tpi = max(T // 8, 1)
cb = torch.randn((T + tpi - 1) // tpi, E, generator=g).cuda()
nsc = b200moe._lib.load().b200moe_router_scratch_bytes(d, E)
wsplit = torch.empty(nsc // 2, dtype=torch.bfloat16, device="cuda")
ops.router_fwd(x, wg, cb, tpi, k, True, wg_split=wsplit)          # fills wsplit
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ts = []
for i in range(iters + 3):
    flush.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    ops.router_fwd(x, None, cb, tpi, k, True, wg_split=wsplit, E=E)
    b.record()
    torch.cuda.synchronize()
    if i >= 3:
        ts.append(a.elapsed_time(b) * 1e3)
ts.sort()
us = ts[len(ts) // 2]
nbytes = T * d * 2 + T * E * 4 + T * k * 8
print(f"router_fwd T={T} d={d} E={E} k={k}: median {us:.1f} us (min {ts[0]:.1f}), {nbytes / us / 1e3:.0f} GB/s")
