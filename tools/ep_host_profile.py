"""cProfile of rank 0's host work in the EP layer step (workload D) under torchrun; prints the top entries."""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402

rank = int(os.environ["RANK"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
w = bench.WORKLOADS["D"]
x_h, go_h, ctx_h = bench.make_inputs(w, seed=1234 + rank)
x, go, ctx = x_h.to(dev), go_h.to(dev), ctx_h.to(dev)
m = bench.build_layer(w, dev, dist.group.WORLD)
m.ep_chunks = int(sys.argv[1]) if len(sys.argv) > 1 else 1
xg = x.clone().requires_grad_(True)


def step():
    m.zero_grad(set_to_none=True)
    xg.grad = None
    y = m(xg, ctx)
    torch.autograd.backward([y, m.aux_loss], [go, torch.ones_like(m.aux_loss)])


for _ in range(5):
    step()
torch.cuda.synchronize()
dist.barrier()
pr = cProfile.Profile()
pr.enable()
for _ in range(20):
    step()
torch.cuda.synchronize()
pr.disable()
if rank == 0:
    pstats.Stats(pr).sort_stats("tottime").print_stats(28)
dist.destroy_process_group()
