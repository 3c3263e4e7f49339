"""Expert-parallel timing breakdown under torchrun: raw all_to_all_single bandwidth at the layer's payload size, and
the EP layer step (workload D) for several ep_chunks settings.  Rank 0 prints one JSON object.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29541 \
        tools/ep_diag.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402


def timed(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / n], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    out = {"world": world}
    w = bench.WORKLOADS["D"]
    rows = w["B"] * w["HW"] * w["k"]
    for frac, name in ((1.0, "a2a_full_ms"), (0.5, "a2a_half_ms")):
        n = int(rows * frac) // world * world
        src = torch.randn(n, w["d"], device=dev).to(torch.bfloat16)
        dst = torch.empty_like(src)
        ms = timed(lambda: dist.all_to_all_single(dst, src))
        sent = n * w["d"] * 2 * (world - 1) / world
        out[name] = {"ms": ms, "bytes_sent_per_rank": sent, "gbs_per_direction": sent / ms / 1e6}
    x_h, go_h, ctx_h = bench.make_inputs(w, seed=1234 + rank)
    x, go, ctx = x_h.to(dev), go_h.to(dev), ctx_h.to(dev)
    m = bench.build_layer(w, dev, dist.group.WORLD)
    xg = x.clone().requires_grad_(True)

    def step():
        m._wcache.clear()
        m.zero_grad(set_to_none=True)
        xg.grad = None
        y = m(xg, ctx)
        torch.autograd.backward([y, m.aux_loss], [go, torch.ones_like(m.aux_loss)])

    for c in (1, 2, 4):
        m.ep_chunks = c
        out[f"step_ms_chunks{c}"] = timed(step, n=10, warm=3)

    def fwd_only():
        with torch.no_grad():
            m(x, ctx)
    m.ep_chunks = 2
    out["fwd_only_ms_chunks2"] = timed(fwd_only)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
