"""Diagnostic for the training-step-contract test: per step and per combine mode, rel-L2 of y and of every gradient
against the oracle (fp16 autocast + GradScaler loop of det_engine.py:44-59, multiscale token counts)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "multimodal-moe_b200"))
import torch  # noqa: E402
from oracle import moe_ref as R  # noqa: E402
import b200moe  # noqa: E402


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-12))


for mode in ("tail", "separate", "epilogue"):
    for use_scaler in (True, False):
        torch.manual_seed(17)
        B, d, F, E, k = 4, 256, 1024, 8, 2
        vs = R.CONTEXT_VOCAB_SIZES
        ref = R.ContextMoERef(d, F, E, k, 1.25, context_vocab_sizes=vs, context_dim=16)
        m = b200moe.ContextMoE(d, F, E, k, 1.25, context_vocab_sizes=vs, context_dim=16)
        m.load_state_dict(ref.state_dict())
        m = m.cuda()
        m.combine = mode
        head = torch.nn.Linear(d, 4).cuda()
        params = list(m.parameters()) + list(head.parameters())
        opt = torch.optim.AdamW(params, lr=1e-4)
        scaler = torch.amp.GradScaler("cuda", enabled=use_scaler)
        g = torch.Generator().manual_seed(3)
        for step, HW in enumerate((225, 400, 625, 400, 225)):
            x = torch.randn(B, HW, d, generator=g)
            c = torch.stack([torch.randint(0, v, (B,), generator=g) for v in vs], 1)
            tgt = torch.randn(B, HW, 4, generator=g)
            ref.load_state_dict(m.state_dict())
            ref.zero_grad(set_to_none=True)
            xg = x.cuda().requires_grad_(True)
            with torch.autocast(device_type="cuda", enabled=use_scaler):
                y = m(xg, c.cuda())
                out = head(y)
            loss = (out.float() - tgt.cuda()).pow(2).mean() + m.aux_loss
            scaler.scale(loss).backward()
            scaler.unscale_(opt)
            xr = x.clone().requires_grad_(True)
            yr = ref(xr, c, logits_value=m.last_aux["logits"].cpu())
            hw_, hb_ = head.weight.detach().cpu().float(), head.bias.detach().cpu().float()
            outr = torch.nn.functional.linear(yr, hw_, hb_)
            ((outr - tgt).pow(2).mean() + ref.aux_loss).backward()
            r = {"y": rel(y, yr), "dx": rel(xg.grad / scaler.get_scale(), xr.grad),
                 "dwg": rel(m.gate.weight.grad, ref.gate.weight.grad)}
            for n in ("w1", "w2", "b1", "b2"):
                r[n] = rel(getattr(m, n).grad, getattr(ref, n).grad)
            # how far the upstream gradient itself is from the oracle's (fp16 head in between)
            print(mode, "scaler" if use_scaler else "fp32", step, HW, " ".join(f"{k_}={v:.2e}" for k_, v in r.items()),
                  "dropped", int(m.last_aux["dropped"]), flush=True)
            torch.nn.utils.clip_grad_norm_(params, 0.1)
            scaler.step(opt)
            scaler.update()
            opt.zero_grad()
