"""Kineto trace of rank 0's EP layer step (workload D) under torchrun -> gpurun_out/ep_trace_rank0.json (chrome trace)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402
import bench  # noqa: E402

rank = int(os.environ["RANK"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
w = bench.WORKLOADS[os.environ.get("EP_TRACE_WORKLOAD", "D")]
x_h, go_h, ctx_h = bench.make_inputs(w, seed=1234 + rank)
x, go, ctx = x_h.to(dev), go_h.to(dev), ctx_h.to(dev)
m = bench.build_layer(w, dev, dist.group.WORLD)
m.ep_chunks = int(sys.argv[1]) if len(sys.argv) > 1 else 1
xg = x.clone().requires_grad_(True)


def step():
    m._wcache.clear()
    m.zero_grad(set_to_none=True)
    xg.grad = None
    y = m(xg, ctx)
    torch.autograd.backward([y, m.aux_loss], [go, torch.ones_like(m.aux_loss)])


for _ in range(5):
    step()
torch.cuda.synchronize()
dist.barrier()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
if rank == 0:
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    prof.export_chrome_trace(os.path.join(ROOT, "gpurun_out", "ep_trace_rank0.json"))
    print("trace written")
dist.destroy_process_group()
