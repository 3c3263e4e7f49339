"""Timeline of ONE graph-replayed fwd+bwd step (what bench.py's `value` times): every kernel's start offset, duration and
stream from torch.profiler (CUPTI), so that the critical path of the latency-bound workloads can be read off.

    python tools/graph_timeline.py [A|C|D] [B] [HW] > gpurun_out/timeline_A.txt
"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "A"
w = dict(bench.WORKLOADS[name])
if len(sys.argv) > 3:
    w.update(B=int(sys.argv[2]), HW=int(sys.argv[3]))
dev = torch.device("cuda", 0)
m = bench.build_layer(w, dev)
x, go, ctx = bench.make_inputs(w, device=dev)
xg = x.clone().requires_grad_(True)
flush = bench.L2Flusher(dev)


def step():
    m._wcache.clear()
    m.zero_grad(set_to_none=True)
    xg.grad = None
    y = m(xg, ctx)
    torch.autograd.backward([y, m.aux_loss], [go, torch.ones_like(m.aux_loss)])


side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for _ in range(3):
        step()
    m.zero_grad(set_to_none=True)
    xg.grad = None
    m.aux_loss, m.last_aux = None, {}
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize()
graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(graph):
    step()
for _ in range(5):
    flush(); graph.replay()
torch.cuda.synchronize()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        flush(); graph.replay()
    torch.cuda.synchronize()
path = os.path.join(tempfile.gettempdir(), "b200moe_timeline.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"]
      if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
ev.sort(key=lambda e: e["ts"])
# split at the L2-flush fills; print the last replay
starts = [i for i, e in enumerate(ev) if "FillFunctor" in e["name"] and e["dur"] > 20]
last = ev[starts[-1] + 1:]
t0 = last[0]["ts"]
end = max(e["ts"] + e["dur"] for e in last)
print(f"workload {name} tokens {w['B'] * w['HW']}: {len(last)} kernels, span {end - t0:.1f} us")
print(f"{'start us':>9} {'dur us':>8} {'end us':>8} {'stream':>6}  kernel")
for e in last:
    nm = e["name"].replace("b200moe::", "").replace("void ", "")
    print(f"{e['ts'] - t0:9.1f} {e['dur']:8.1f} {e['ts'] - t0 + e['dur']:8.1f} {str(e['args'].get('stream')):>6}  {nm[:90]}")
