"""Expert-parallel parity check, run under torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/ep_check.py

Every rank runs the EP layer on its own tokens; every rank also runs the single-GPU layer (all experts local, same
seed => same weights) on the concatenated, rank-major batch.  Checked: routing and drops bit-exact against the
single-GPU run, outputs / dx / expert gradients / router gradient within the bf16 tolerance.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "multimodal-moe_b200"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


def varying_tokens_and_overflow(rank, world, dev):
    """Multiscale training (dataloader.yml:22): the token count changes per step AND per rank.  With ep_max_tokens
    declared the peer-memory arena serves all of them; without it, a batch whose rows do not fit raises on EVERY rank
    (at the next backward) instead of on one rank while the others hang."""
    import b200moe
    d, F, E, k = 128, 256, 2 * world, 2
    torch.manual_seed(1)
    m = b200moe.ContextMoE(d, F, E, k, 1.0, ep_group=dist.group.WORLD, ep_p2p=None, ep_max_tokens=2 * 625).to(dev)
    torch.manual_seed(1)
    m1 = b200moe.ContextMoE(d, F, E, k, 1.0).to(dev)
    ok = True
    for step, hw_by_rank in enumerate([(225, 400), (625, 225), (400, 625), (625, 625)]):
        hws = [hw_by_rank[r % 2] for r in range(world)]
        g = torch.Generator().manual_seed(50 + step)
        xs = [torch.randn(2, hw, d, generator=g).to(torch.bfloat16).to(dev) for hw in hws]
        x = xs[rank].clone().requires_grad_(True)
        y = m(x)
        y.float().sum().backward()
        # reference: every rank's tokens through the single-GPU layer (ragged batch: flat tokens, one "image" each)
        xa = torch.cat([t.reshape(-1, d) for t in xs]).clone().requires_grad_(True)
        y1 = m1(xa)
        y1.float().sum().backward()
        lo = sum(2 * h for h in hws[:rank])
        same = torch.equal(y.reshape(-1, d), y1[lo:lo + 2 * hws[rank]]) and \
            torch.equal(x.grad.reshape(-1, d), xa.grad[lo:lo + 2 * hws[rank]])
        if not same:
            print(f"[rank {rank}] FAILED varying tokens step {step} hws={hws}", flush=True)
        ok = ok and same
    path = "p2p" if "_p2p_arena" in m.__dict__ else "nccl"
    raised = None
    if path == "p2p":
        torch.manual_seed(2)
        small = b200moe.ContextMoE(d, F, E, k, 1.0, ep_group=dist.group.WORLD).to(dev)
        with torch.no_grad():
            small(torch.randn(1, 128, d, device=dev).to(torch.bfloat16))        # arena sized for 128 tokens per rank
        xb = torch.randn(4, 512, d, device=dev).to(torch.bfloat16).requires_grad_(True)
        try:
            ys = small(xb)
            ar = small.__dict__.get("_p2p_arena")
            torch.cuda.synchronize()
            print(f"[rank {rank}] overflow case: arena rows_cap={ar.rows_cap} T_cap={ar.T_cap} pending={ar.pending} "
                  f"flag={ar.flag_host.tolist()} event={ar.flag_event is not None}", flush=True)
            ys.float().sum().backward()
            raised = False
        except RuntimeError as e:
            raised = "arena overflow" in str(e)
            if not raised:
                print(f"[rank {rank}] unexpected error: {e}", flush=True)
        ok = ok and bool(raised)
    print(f"[rank {rank}] varying tokens per step and rank over {path}: {'OK' if ok else 'FAILED'}"
          f" (overflow raised on this rank: {raised})", flush=True)
    return ok


def ddp_gradient_scale(rank, world, dev):
    """The DDP rule (b200moe/checkpoint.py): under real DistributedDataParallel with the experts ignored, every gradient
    -- replicated (DDP mean) and sharded (expert_grad_scale = 1/W) -- equals 1/W times the gradient of the single-GPU
    layer on the concatenated batch, i.e. the gradient of the MEAN of the per-rank losses; and the EP-aware clip leaves
    the replicated parameters bit-identical across the ranks."""
    import b200moe
    from torch.nn.parallel import DistributedDataParallel as DDP
    B, HW, d, F, E, k = 2, 200, 256, 512, 2 * world, 2
    torch.manual_seed(9)
    m = b200moe.ContextMoE(d, F, E, k, 1.25, context_vocab_sizes=(6, 10), ep_group=dist.group.WORLD).to(dev)
    torch.manual_seed(9)
    m1 = b200moe.ContextMoE(d, F, E, k, 1.25, context_vocab_sizes=(6, 10)).to(dev)
    b200moe.ddp_ignore_experts(m)
    ddp = DDP(m, device_ids=[dev.index])
    g = torch.Generator().manual_seed(77)
    x_all = torch.randn(world * B, HW, d, generator=g).to(torch.bfloat16).to(dev)
    go_all = torch.randn(world * B, HW, d, generator=g).to(torch.bfloat16).to(dev)
    c_all = torch.stack([torch.randint(0, v, (world * B,), generator=g) for v in (6, 10)], 1).to(dev)
    sl = slice(rank * B, (rank + 1) * B)
    y = ddp(x_all[sl], c_all[sl])
    y.backward(go_all[sl])
    y1 = m1(x_all, c_all)
    y1.backward(go_all)
    EL = E // world
    esl = slice(rank * EL, (rank + 1) * EL)
    inv = 1.0 / world
    checks = {"gate (DDP mean)": rel(m.gate.weight.grad, m1.gate.weight.grad * inv),
              "ctx_proj (DDP mean)": rel(m.ctx_proj.weight.grad, m1.ctx_proj.weight.grad * inv),
              "w1 (expert_grad_scale)": rel(m.w1.grad, m1.w1.grad[esl] * inv),
              "b2 (expert_grad_scale)": rel(m.b2.grad, m1.b2.grad[esl] * inv)}
    ok = all(v < 1e-2 for v in checks.values())
    norm = b200moe.clip_grad_norm_(ddp, 0.1)
    ref_norm = torch.sqrt(sum((p.grad.float() * inv).pow(2).sum() for p in m1.parameters() if p.grad is not None))
    ok = ok and abs(float(norm) - float(ref_norm)) <= 1e-2 * float(ref_norm)
    parts = [torch.empty_like(m.gate.weight.grad) for _ in range(world)]
    dist.all_gather(parts, m.gate.weight.grad)
    ok = ok and all(torch.equal(parts[0], q) for q in parts[1:])
    print(f"[rank {rank}] DDP + expert parallelism: " + " ".join(f"{n}={v:.2e}" for n, v in checks.items()) +
          f" clip norm {float(norm):.4f} vs {float(ref_norm):.4f}: {'OK' if ok else 'FAILED'}", flush=True)
    return ok


def main():
    import b200moe
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    cases = [(2, 300, 256, 512, 8, 2, 1.0), (1, 777, 512, 1024, 16, 2, 1.25), (2, 64, 128, 256, 2 * world, 1, 0.5),
             (4, 400, 256, 1024, 8, 2, None)]
    # every case over peer memory (the default where available; cf=None has no row bound and uses NCCL) and over NCCL
    for (B, HW, d, F, E, k, cf), p2p in [(c, m) for c in cases for m in (None, False)]:
        vs = (6, 10, 4, 6, 4)
        torch.manual_seed(7)
        m_ep = b200moe.ContextMoE(d, F, E, k, cf, context_vocab_sizes=vs, ep_group=dist.group.WORLD,
                                  ep_p2p=p2p).to(dev)
        torch.manual_seed(7)
        m_1 = b200moe.ContextMoE(d, F, E, k, cf, context_vocab_sizes=vs).to(dev)
        with torch.no_grad():                                       # skew the router so that capacity bites
            m_ep.gate.weight[0] *= 3.0
            m_1.gate.weight[0] *= 3.0
        g = torch.Generator().manual_seed(1000)
        x_all = torch.randn(world * B, HW, d, generator=g).to(torch.bfloat16).to(dev)
        go_all = torch.randn(world * B, HW, d, generator=g).to(torch.bfloat16).to(dev)
        c_all = torch.stack([torch.randint(0, v, (world * B,), generator=g) for v in vs], 1).to(dev)
        sl = slice(rank * B, (rank + 1) * B)

        for _ in range(2):                                          # twice: the second pass reuses the peer arena
            m_ep.zero_grad(set_to_none=True)
            x = x_all[sl].clone().requires_grad_(True)
            y = m_ep(x, c_all[sl])
            y.backward(go_all[sl])
        a_ep = dict(m_ep.last_aux)
        path = "p2p" if "_p2p_arena" in m_ep.__dict__ else "nccl"
        if path == "p2p":
            # a second forward before the first backward must not touch the arena the pending backward reads
            xa = x_all[sl].clone().requires_grad_(True)
            ya = m_ep(xa, c_all[sl])
            with torch.no_grad():
                yb = m_ep(x_all[sl], c_all[sl])                     # eval-style call in between (NCCL path)
            m_ep.zero_grad(set_to_none=True)
            ya.backward(go_all[sl])
            ok = ok and torch.equal(ya, y) and torch.equal(yb, y) and torch.equal(xa.grad, x.grad)
            if not (torch.equal(ya, y) and torch.equal(yb, y) and torch.equal(xa.grad, x.grad)):
                print(f"[rank {rank}] FAILED interleaved forward on the p2p path", flush=True)

        x1 = x_all.clone().requires_grad_(True)
        y1 = m_1(x1, c_all)
        y1.backward(go_all)
        a_1 = dict(m_1.last_aux)
        torch.cuda.synchronize()

        T = B * HW
        tsl = slice(rank * T, (rank + 1) * T)
        EL = E // world
        esl = slice(rank * EL, (rank + 1) * EL)
        checks = {
            "topk_idx bit-exact": torch.equal(a_ep["topk_idx"], a_1["topk_idx"][tsl]),
            "drops bit-exact": torch.equal(a_ep["row_of_assignment"] >= 0, a_1["row_of_assignment"][tsl] >= 0),
            "y": rel(y, y1[sl]) < 1e-2,
            "dx": rel(x.grad, x1.grad[sl]) < 1e-2,
            "dw1": rel(m_ep.w1.grad, m_1.w1.grad[esl]) < 1e-2,
            "dw2": rel(m_ep.w2.grad, m_1.w2.grad[esl]) < 1e-2,
            "db1": rel(m_ep.b1.grad, m_1.b1.grad[esl]) < 1e-2,
            "db2": rel(m_ep.b2.grad, m_1.b2.grad[esl]) < 1e-2,
        }
        dwg = m_ep.gate.weight.grad.clone()
        dist.all_reduce(dwg)
        checks["dwg (sum over ranks)"] = rel(dwg, m_1.gate.weight.grad) < 1e-2
        dcp = m_ep.ctx_proj.weight.grad.clone()
        dist.all_reduce(dcp)
        checks["dctx_proj (sum over ranks)"] = rel(dcp, m_1.ctx_proj.weight.grad) < 1e-2
        bad = [n for n, v in checks.items() if not v]
        print(f"[rank {rank}] {path} B={B} HW={HW} d={d} F={F} E={E} k={k} cf={cf}: dropped(ep local)={int(a_ep['dropped'])} "
              f"dropped(1gpu)={int(a_1['dropped'])} rel_y={rel(y, y1[sl]):.2e} rel_dx={rel(x.grad, x1.grad[sl]):.2e} "
              f"{'OK' if not bad else 'FAILED ' + str(bad)}", flush=True)
        ok = ok and not bad
    ok = varying_tokens_and_overflow(rank, world, dev) and ok
    ok = ddp_gradient_scale(rank, world, dev) and ok
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.destroy_process_group()
    if int(flag) != 0:
        raise SystemExit(1)
    if rank == 0:
        print("EP_CHECK_OK", flush=True)


if __name__ == "__main__":
    main()
