"""unpermute_bwd alone:  python tools/unp_bench.py T d E  (B200MOE_UNP_TT=2|4|8 overrides the tokens-per-warp variant)"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "multimodal-moe_b200"))
import torch
import b200moe
from b200moe import ops
T, d, E = (int(a) for a in sys.argv[1:4])
k = 2
b200moe.load_library()
g = torch.Generator(device="cuda").manual_seed(0)
R = T * k
dxp = torch.randn(R, d, device="cuda", generator=g).to(torch.bfloat16)
roa = torch.randperm(R, device="cuda", generator=g).int().view(T, k)
dl = torch.randn(T, E, device="cuda", generator=g) * 0.01
wg = torch.randn(E, d, device="cuda", generator=g)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ts = []
for i in range(10):
    flush.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); ops.unpermute_bwd(dxp, roa, dl, wg, T, k, d); b.record()
    torch.cuda.synchronize()
    if i >= 2:
        ts.append(a.elapsed_time(b) * 1e3)
ts.sort()
us = ts[len(ts) // 2]
nb = 2 * R * d + 2 * T * d
print(f"unpermute_bwd T={T} d={d} E={E} TT={os.environ.get('B200MOE_UNP_TT', 'auto')}: {us:.1f} us  {nb / us / 1e3:.0f} GB/s", flush=True)
