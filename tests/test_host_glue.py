"""CPU: host-side glue around the layer — context vocabulary and wire-through against
the reference's exported fields, expert-parallel state dicts over gloo (world size 2), DDP ignore list."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

from oracle import moe_ref as R


def test_vocabulary_matches_oracle_and_reference_lists():
    from b200moe import context as cx
    assert cx.CONTEXT_FIELDS == R.CONTEXT_FIELDS
    assert cx.CONTEXT_VOCAB == R.CONTEXT_VOCAB                  # the oracle pins these against the reference csv (K4)
    assert cx.CONTEXT_VOCAB_SIZES == (6, 10, 4, 6, 4)
    assert all(v[0] == "missing" for v in cx.CONTEXT_VOCAB.values())


def test_context_ids_from_coco_image_records():
    from b200moe import context as cx
    rec = {"file_name": "000001.jpg", "solar_context_bin": "low_sun(0..15)", "scraped_weather": "rain",
           "time_of_day": "day", "road_type": "highway", "road_condition": "wet"}
    ids = cx.context_ids(rec)
    assert ids == [3, cx.CONTEXT_VOCAB["scraped_weather"].index("rain"), cx.CONTEXT_VOCAB["time_of_day"].index("day"),
                   cx.CONTEXT_VOCAB["road_type"].index("highway"), cx.CONTEXT_VOCAB["road_condition"].index("wet")]
    assert cx.context_ids({"solar_context_bin": None, "scraped_weather": float("nan"), "road_type": "moon"}) == [0] * 5
    t = cx.attach_context({"boxes": torch.zeros(0, 4)}, rec)
    assert t["context"].dtype == torch.long and t["context"].tolist() == ids
    batch = cx.batch_context([t, cx.attach_context({}, {})])
    assert batch.shape == (2, 5) and batch[1].tolist() == [0] * 5
    assert cx.batch_context([{"boxes": 1}]) is None and cx.batch_context(None) is None


def test_context_hook_reaches_layers():
    from b200moe import context as cx

    class Layer(nn.Module):
        def __init__(self):
            super().__init__()
            self.seen = "unset"

        def set_context(self, c):
            self.seen = c

    class Det(nn.Module):                                       # RTDETR.forward(x, targets=None) signature, rtdetr.py:32
        def __init__(self):
            super().__init__()
            self.enc = nn.ModuleList([Layer(), Layer()])

        def forward(self, x, targets=None):
            return x

    det = Det()
    h = cx.install_context_hook(det)
    tg = [cx.attach_context({}, {"time_of_day": "night"}), cx.attach_context({}, {"road_condition": "snow"})]
    det(torch.zeros(2, 3), targets=tg)
    assert all(l_.seen.shape == (2, 5) for l_ in det.enc) and det.enc[0].seen[0, 2] == 2
    det(torch.zeros(2, 3), tg)                                   # positional targets
    assert det.enc[1].seen[1, 4] == cx.CONTEXT_VOCAB["road_condition"].index("snow")
    det(torch.zeros(2, 3))
    assert all(l_.seen is None for l_ in det.enc)
    h.remove()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _ckpt_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import b200moe
        from b200moe import checkpoint as ck
        torch.manual_seed(5)
        single = b200moe.ContextMoE(128, 256, 4, 2, context_vocab_sizes=(6, 10))
        torch.manual_seed(5)
        sharded = b200moe.ContextMoE(128, 256, 4, 2, context_vocab_sizes=(6, 10), ep_group=dist.group.WORLD)
        wrap = nn.Sequential(nn.Linear(4, 4), sharded)
        assert sharded.w1.shape[0] == 2 and sharded.ep_rank == rank
        # same seed => the sharded model is a slice of the single-GPU model
        assert torch.equal(sharded.w1, single.w1[rank * 2:(rank + 1) * 2])
        assert ck.expert_param_names(wrap) == ["1.w1", "1.b1", "1.w2", "1.b2"]
        assert ck.ddp_ignore_experts(wrap) == ["1.w1", "1.b1", "1.w2", "1.b2"]
        assert wrap._ddp_params_and_buffers_to_ignore == ["1.w1", "1.b1", "1.w2", "1.b2"]
        full = ck.full_state_dict(wrap)
        for p in ck.EXPERT_PARAMS:
            assert torch.equal(full[f"1.{p}"], getattr(single, p).detach())
        assert torch.equal(full["1.gate.weight"], single.gate.weight.detach())
        # round trip into a differently initialised sharded model, and into a single-GPU model
        torch.manual_seed(99)
        other = nn.Sequential(nn.Linear(4, 4), b200moe.ContextMoE(128, 256, 4, 2, context_vocab_sizes=(6, 10),
                                                                 ep_group=dist.group.WORLD))
        ck.load_full_state_dict(other, full)
        assert torch.equal(other[1].w2, sharded.w2) and torch.equal(other[1].gate.weight, sharded.gate.weight)
        torch.manual_seed(98)
        one = b200moe.ContextMoE(128, 256, 4, 2, context_vocab_sizes=(6, 10))
        one.load_state_dict({k[2:]: v for k, v in full.items() if k.startswith("1.")})
        assert torch.equal(one.b1, single.b1)
        if rank == 0:
            out.put("ok")
    finally:
        dist.destroy_process_group()


def test_ep_state_dict_round_trip_world2():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_ckpt_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
    assert all(p.exitcode == 0 for p in procs)
    assert out.get(timeout=5) == "ok"


def _clip_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import b200moe
        from b200moe import checkpoint as ck
        torch.manual_seed(3)
        moe = b200moe.ContextMoE(128, 128, 4, 2, ep_group=dist.group.WORLD)
        model = nn.Sequential(nn.Linear(8, 8), moe)
        assert moe.expert_grad_scale == 1.0
        ck.ddp_ignore_experts(model)
        assert moe.expert_grad_scale == 0.5                       # DDP's mean-over-ranks convention (2 ranks)
        g = torch.Generator().manual_seed(11)                     # the same "global" gradient on both ranks ...
        sharded_names = set(ck.expert_param_names(model))
        rep = {n: torch.randn(p.shape, generator=g) for n, p in model.named_parameters() if n not in sharded_names}
        full = {p: torch.randn((4,) + tuple(getattr(moe, p).shape[1:]), generator=g) for p in ck.EXPERT_PARAMS}
        for n, p in model.named_parameters():
            if n in sharded_names:
                p.grad = full[n.split(".")[-1]][rank * 2:(rank + 1) * 2].clone()   # ... each rank holds its expert shard
            else:
                p.grad = rep[n].clone()
        ref_sq = sum(float(v.pow(2).sum()) for v in rep.values()) + sum(float(v.pow(2).sum()) for v in full.values())
        norm = ck.clip_grad_norm_(model, 0.1)
        assert abs(float(norm) - ref_sq ** 0.5) < 1e-3 * ref_sq ** 0.5
        coef = 0.1 / (ref_sq ** 0.5 + 1e-6)
        assert torch.allclose(model[0].weight.grad, rep["0.weight"] * coef, rtol=1e-5, atol=1e-8)
        assert torch.allclose(moe.w1.grad, full["w1"][rank * 2:(rank + 1) * 2] * coef, rtol=1e-5, atol=1e-8)
        # replicated gradients stay bit-identical across the ranks after the clip
        gathered = [torch.empty_like(moe.gate.weight.grad) for _ in range(world)]
        dist.all_gather(gathered, moe.gate.weight.grad)
        assert torch.equal(gathered[0], gathered[1])
        # torch's own clip would have used a rank-local norm: different coefficients on the two ranks
        local_sq = sum(float(p.grad.pow(2).sum()) for p in model.parameters())
        both = [None, None]
        dist.all_gather_object(both, local_sq)
        assert abs(both[0] - both[1]) > 1e-12

        # EPGradScaler: an overflow in ONE rank's shard is seen by every rank
        opt = torch.optim.SGD(model.parameters(), lr=0.1)
        scaler = ck.EPGradScaler(ep_group=dist.group.WORLD, device="cpu", init_scale=4.0)
        before = model[0].weight.detach().clone()
        for p in model.parameters():
            p.grad = torch.ones_like(p)
        scaler.scale(torch.zeros((), requires_grad=True))          # initialises the scale tensor lazily
        if rank == 1:
            moe.w2.grad[0, 0, 0] = float("inf")
        scaler.unscale_(opt)
        scaler.step(opt)
        scaler.update()
        assert torch.equal(model[0].weight.detach(), before)       # skipped on BOTH ranks
        assert scaler.get_scale() == 2.0                           # and the scale backed off on both
        if rank == 0:
            out.put("ok")
    finally:
        dist.destroy_process_group()


def test_ep_clip_and_scaler_world2():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_clip_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    assert all(p.exitcode == 0 for p in procs)
    assert out.get(timeout=5) == "ok"
