"""Which host op launches the small fill kernels seen between forward and backward of a step?  (torch.profiler, CPU ops
with Python stacks)   python tools/find_fills.py > gpurun_out/find_fills.txt"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402

dev = torch.device("cuda", 0)
w = dict(bench.WORKLOADS["A"])
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


for _ in range(3):
    step()
torch.cuda.synchronize()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CPU, torch.profiler.ProfilerActivity.CUDA],
                            with_stack=True) as prof:
    step()
    torch.cuda.synchronize()
for e in prof.events():
    if e.name in ("aten::fill_", "aten::zero_", "aten::zeros", "aten::ones", "aten::zeros_like", "aten::ones_like",
                  "aten::full", "aten::new_zeros"):
        print(e.name, [str(s) for s in (e.input_shapes or [])], "parent:", e.cpu_parent.name if e.cpu_parent else None)
        for s in (e.stack or [])[:8]:
            print("     ", s)
