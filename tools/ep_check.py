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
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.destroy_process_group()
    if int(flag) != 0:
        raise SystemExit(1)
    if rank == 0:
        print("EP_CHECK_OK", flush=True)


if __name__ == "__main__":
    main()
