"""Diagnostic for the tcgen05 grouped GEMM stages: per-stage relative error and, when a stage is wrong, a coarse
error map of its first 128x128 tile (rows x cols in 8x8 blocks) to tell descriptor / swizzle mistakes apart.

    python tools/gemm_diag.py [out_file]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-moe_b200"))
import b200moe  # noqa: E402
from b200moe import ops  # noqa: E402

out = open(sys.argv[1], "w") if len(sys.argv) > 1 else sys.stdout


def say(*a):
    print(*a, file=out, flush=True)
    if out is not sys.stdout:
        print(*a, flush=True)


def emap(got, ref):
    err = (got.float() - ref.float()).abs()[:128, :128]
    scale = float(ref.float().abs().max()) + 1e-9
    rows = []
    for i in range(0, 128, 8):
        row = ""
        for j in range(0, 128, 8):
            v = float(err[i:i + 8, j:j + 8].max()) / scale
            row += "." if v < 0.02 else ("o" if v < 0.2 else "X")
        rows.append(row)
    return "\n".join(rows)


def rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


def layout(E, n, extra=128):
    offs = torch.arange(E + 1, dtype=torch.int32) * n
    R = E * n
    rows_cap = R + extra
    meta = torch.zeros(8, dtype=torch.int32)
    meta[0], meta[1] = R, R // 128
    te = torch.full((rows_cap // 128,), -1, dtype=torch.int32)
    for e in range(E):
        te[e * n // 128:(e + 1) * n // 128] = e
    return offs.cuda(), meta.cuda(), te.cuda(), R, rows_cap


def main():
    b200moe.load_library()
    torch.manual_seed(0)
    say("device", torch.cuda.get_device_name(0))
    for (E, d, F, n) in [(2, 128, 256, 256), (2, 256, 1024, 512), (2, 1024, 4096, 42 * 128)]:
        offs, meta, te, R, rows_cap = layout(E, n)
        say(f"=== E={E} d={d} F={F} rows/expert={n}")
        xp = torch.randn(rows_cap, d, device="cuda").to(torch.bfloat16)
        h = torch.randn(rows_cap, F, device="cuda").to(torch.bfloat16)
        w1 = (torch.randn(E, F, d, device="cuda") * d ** -0.5).to(torch.bfloat16)
        w2 = (torch.randn(E, d, F, device="cuda") * F ** -0.5).to(torch.bfloat16)
        z1 = torch.zeros(E, F, device="cuda")
        z2 = torch.zeros(E, d, device="cuda")
        stages = []
        try:
            pre, _ = ops.ffn_up(xp, w1, z1, meta, te)
            torch.cuda.synchronize()
            ref = torch.cat([xp[e * n:(e + 1) * n].float() @ w1[e].float().t() for e in range(E)])
            stages.append(("up   (A K-major, B K-major)", pre[:R], ref))
            yo = ops.ffn_down(h, w2, z2, meta, te)
            torch.cuda.synchronize()
            ref = torch.cat([h[e * n:(e + 1) * n].float() @ w2[e].float().t() for e in range(E)])
            stages.append(("down (A K-major, B K-major)", yo[:R], ref))
            dxp = ops.ffn_dgrad_up(h, w1, meta, te)
            torch.cuda.synchronize()
            ref = torch.cat([h[e * n:(e + 1) * n].float() @ w1[e].float() for e in range(E)])
            stages.append(("dgrad_up (A K-major, B MN-major)", dxp[:R], ref))
            dw = ops.ffn_wgrad(xp, h, offs, E)
            torch.cuda.synchronize()
            ref = torch.stack([xp[e * n:(e + 1) * n].float().t() @ h[e * n:(e + 1) * n].float() for e in range(E)])
            stages.append(("wgrad (A MN-major, B MN-major)", dw.reshape(E * d, F), ref.reshape(E * d, F)))
        except Exception as ex:  # noqa: BLE001
            say("EXCEPTION", repr(ex))
        for name, got, ref in stages:
            r = rel(got, ref)
            say(f"{name:40s} rel={r:.3e} {'OK' if r < 6e-3 else 'WRONG'}")
            if r >= 6e-3:
                say(emap(got, ref))
                say("got[0,:8]", got[0, :8].float().tolist())
                say("ref[0,:8]", ref[0, :8].float().tolist())


if __name__ == "__main__":
    main()
