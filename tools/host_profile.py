"""Host-side cost of one end-to-end step (pinned H2D, fwd+bwd through the public module, loss read-back):
python tools/host_profile.py [A|C|D]  -> cProfile top entries.  Run on the GPU box."""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402

w = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "A"]
dev = torch.device("cuda", 0)
m = bench.build_layer(w, dev)
x_h, go_h, ctx_h = bench.make_inputs(w)
go = go_h.to(dev)
xp_h, ctxp_h = x_h.pin_memory(), ctx_h.pin_memory()


def step():
    m.zero_grad(set_to_none=True)
    xd = xp_h.to(dev, non_blocking=True).requires_grad_(True)
    cd = ctxp_h.to(dev, non_blocking=True)
    y = m(xd, cd)
    loss = (y.float() * go).sum() + m.aux_loss
    loss.backward()
    return float(loss.item())


for _ in range(20):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(30)
