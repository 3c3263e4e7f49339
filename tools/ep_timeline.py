"""Kernel timeline of ONE expert-parallel fwd+bwd step (workload D, eager launches, peer-memory transport) on rank 0:
start offset, duration and stream of every kernel / memcpy (torch.profiler, CUPTI), so that barrier waits, the NCCL
all-gather and launch gaps can be told apart from kernel time.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29547 \
        tools/ep_timeline.py > gpurun_out/ep_timeline.txt
"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
w = dict(bench.WORKLOADS["D"])
m = bench.build_layer(w, dev, ep_group=dist.group.WORLD)
x, go, ctx = bench.make_inputs(w, seed=1234 + rank, device=dev)
xg = x.clone().requires_grad_(True)
flush = bench.L2Flusher(dev)


def step():
    m._wcache.clear()
    m.zero_grad(set_to_none=True)
    xg.grad = None
    y = m(xg, ctx)
    torch.autograd.backward([y, m.aux_loss], [go, torch.ones_like(m.aux_loss)])


for _ in range(5):
    flush(); step()
torch.cuda.synchronize()
dist.barrier()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        flush(); step()
    torch.cuda.synchronize()
dist.barrier()
if rank == 0:
    path = os.path.join(tempfile.gettempdir(), "b200moe_ep_timeline.json")
    prof.export_chrome_trace(path)
    ev = [e for e in json.load(open(path))["traceEvents"]
          if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    ev.sort(key=lambda e: e["ts"])
    starts = [i for i, e in enumerate(ev) if "FillFunctor" in e["name"] and e["dur"] > 20]
    last = ev[starts[-1] + 1:]
    t0 = last[0]["ts"]
    end = max(e["ts"] + e["dur"] for e in last)
    print(f"workload D, {world} ranks, {w['B'] * w['HW']} tokens per rank: {len(last)} kernels, span {end - t0:.1f} us")
    print(f"{'start us':>9} {'dur us':>8} {'end us':>8} {'gap us':>7} {'stream':>6}  kernel")
    prev_end = {}
    for e in last:
        nm = e["name"].replace("b200moe::", "").replace("void ", "")
        s = e["args"].get("stream")
        gap = e["ts"] - prev_end.get(s, e["ts"])
        prev_end[s] = e["ts"] + e["dur"]
        print(f"{e['ts'] - t0:9.1f} {e['dur']:8.1f} {e['ts'] - t0 + e['dur']:8.1f} {gap:7.1f} {str(s):>6}  {nm[:80]}")
dist.destroy_process_group()
