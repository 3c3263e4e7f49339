"""One eager fwd+bwd step of a bench workload, for ncu:  python tools/profile_step.py <A|C|D> <B> <HW> [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402

name, B, HW = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
w = dict(bench.WORKLOADS[name], B=B, HW=HW)
dev = torch.device("cuda", 0)
m = bench.build_layer(w, dev)
x, go, ctx = bench.make_inputs(w, device=dev)
xg = x.clone().requires_grad_(True)
for _ in range(steps):
    m._wcache.clear()
    m.zero_grad(set_to_none=True)
    xg.grad = None
    y = m(xg, ctx)
    torch.autograd.backward([y, m.aux_loss], [go, torch.ones_like(m.aux_loss)])
torch.cuda.synchronize()
print("ok", name, B * HW, "kept", int(m.last_aux["meta"][3]))
