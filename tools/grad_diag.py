"""Where does the router-gradient error come from?  For the shapes of tests/test_gpu_layer.py, prints the rel-L2 of
every gradient against the fp32 oracle and splits the gate-gradient chain into its stages:

    dy (bf16)  ->  dgate_row = <dy, yo>  ->  dlogits (A.12, fp32)  ->  dWg = dlogits^T x

by feeding each kernel stage the oracle's value of its input (run on the GPU box: python tools/grad_diag.py)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "multimodal-moe_b200"))

import torch  # noqa: E402

from oracle import moe_ref as R  # noqa: E402


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-12))


def main():
    import b200moe
    from b200moe import ops
    cases = [(8, 400, 256, 1024, 8, 2, None, False), (8, 400, 256, 1024, 8, 2, 1.25, True),
             (4, 230, 512, 2048, 16, 2, 1.25, True), (2, 333, 128, 256, 4, 1, 0.5, False),
             (1, 200, 1024, 4096, 64, 2, 2.0, False)]
    for B, HW, d, F, E, k, cf, ctx in cases:
        torch.manual_seed(B * HW)
        vs = R.CONTEXT_VOCAB_SIZES if ctx else ()
        ref = R.ContextMoERef(d, F, E, k, cf, context_vocab_sizes=vs, context_dim=16)
        with torch.no_grad():
            for p in (ref.w1, ref.w2):
                p.copy_(p.to(torch.bfloat16).float())
        m = b200moe.ContextMoE(d, F, E, k, cf, context_vocab_sizes=vs, context_dim=16)
        m.load_state_dict(ref.state_dict())
        m = m.cuda()
        g = torch.Generator().manual_seed(1234)
        x = torch.randn(B, HW, d, generator=g).to(torch.bfloat16)
        go = torch.randn(B, HW, d, generator=g)
        c = torch.stack([torch.randint(0, v, (B,), generator=g) for v in R.CONTEXT_VOCAB_SIZES], 1) if ctx else None
        for go_bf16 in (False, True):
            gof = go.to(torch.bfloat16).float() if go_bf16 else go
            m.zero_grad(set_to_none=True)
            ref.zero_grad(set_to_none=True)
            xg = x.cuda().requires_grad_(True)
            y = m(xg, None if c is None else c.cuda())
            ((y.float() * gof.cuda()).sum() + m.aux_loss).backward()
            torch.cuda.synchronize()
            xr = x.float().clone().requires_grad_(True)
            # the oracle's logits as a leaf so that dL/dlogits can be read back
            lgk = m.last_aux["logits"].cpu()
            x2 = xr.reshape(-1, d)
            lg = ref.logits(x2, ref.context_bias(c), HW)
            lg = lg + (lgk - lg).detach()
            lg.retain_grad()
            yr = ref(xr, c, logits_override=lg)
            ((yr * gof).sum() + ref.aux_loss).backward()
            out = {"y": rel(y, yr), "dx": rel(xg.grad, xr.grad), "dwg": rel(m.gate.weight.grad, ref.gate.weight.grad)}
            for n in ("w1", "b1", "w2", "b2"):
                out["d" + n] = rel(getattr(m, n).grad, getattr(ref, n).grad)
            if ctx:
                out["dctx_proj"] = rel(m.ctx_proj.weight.grad, ref.ctx_proj.weight.grad)
                out["dctx_emb0"] = rel(m.ctx_embed[0].weight.grad, ref.ctx_embed[0].weight.grad)
            # stage split
            pl, ws = m._last_ws
            v = lambda n: pl.view(ws, n)  # noqa: E731
            T = B * HW
            dyb = gof.reshape(T, d).to(torch.bfloat16).cuda().contiguous()
            _, dgate_k = ops.combine_bwd(dyb, v("yo"), v("token_of_row"), v("gate_of_row"), v("meta"), pl.rows_cap)
            # oracle dgate per kept row: <go[t], FFN_e(x[t])>
            tor = v("token_of_row").cpu().long()
            roa = v("row_of_assignment").cpu().long()
            idx = v("topk_idx").cpu().long()
            dgate_o = torch.zeros(pl.rows_cap)
            with torch.no_grad():
                xf = x.float().reshape(T, d)
                gf = gof.reshape(T, d)
                for e in range(E):
                    for s in range(k):
                        sel = torch.nonzero((idx[:, s] == e) & (roa[:, s] >= 0)).squeeze(1)
                        if sel.numel():
                            dgate_o[roa[sel, s]] = (ref.experts(xf[sel], e) * gf[sel]).sum(-1)
            kept = tor >= 0
            out["dgate"] = rel(dgate_k.cpu()[kept], dgate_o[kept])
            daux = torch.ones((), device="cuda")
            tpe = v("tokens_per_expert")
            dl_kk = ops.router_bwd_dlogits(v("logits"), v("topk_idx"), v("row_of_assignment"), dgate_k, tpe, daux,
                                           True, m.aux_loss_weight, m.z_loss_weight)
            dl_ko = ops.router_bwd_dlogits(v("logits"), v("topk_idx"), v("row_of_assignment"), dgate_o.cuda(), tpe, daux,
                                           True, m.aux_loss_weight, m.z_loss_weight)
            out["dlogits(kernel dgate)"] = rel(dl_kk, lg.grad)
            out["dlogits(oracle dgate)"] = rel(dl_ko, lg.grad)
            dwg_o = ops.router_bwd_dwg(lg.grad.cuda().contiguous(), x.reshape(T, d).cuda().contiguous())
            out["dwg(oracle dlogits)"] = rel(dwg_o, ref.gate.weight.grad)
            # magnitude of the three dlogits sources
            print(f"B={B} HW={HW} d={d} F={F} E={E} k={k} cf={cf} ctx={ctx} go_bf16={go_bf16}: " +
                  " ".join(f"{n}={val:.2e}" for n, val in out.items()), flush=True)


if __name__ == "__main__":
    main()
