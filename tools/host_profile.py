"""cProfile of the eager ContextMoE step (host time per step is what bounds the eager module at small sizes):
    python tools/host_profile.py [A|C|D] > gpurun_out/host_profile.txt"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402

dev = torch.device("cuda", 0)
w = dict(bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "A"])
m = bench.build_layer(w, dev)
x, go, ctx = bench.make_inputs(w, device=dev)
xg = x.clone().requires_grad_(True)
one = torch.ones((), device=dev)


def step():
    m._wcache.clear()
    m.zero_grad(set_to_none=True)
    xg.grad = None
    y = m(xg, ctx)
    torch.autograd.backward([y, m.aux_loss], [go, one])


for _ in range(20):
    step()
torch.cuda.synchronize()
N = 300
t0 = time.perf_counter()
for _ in range(N):
    step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host time per step {1e6 * (t1 - t0) / N:.1f} us (queue drained after {1e6 * (t2 - t1):.0f} us more)")
pr = cProfile.Profile()
pr.enable()
for _ in range(N):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
