"""GPU parity: each tcgen05 grouped GEMM stage against a plain fp32 torch contraction of the same bf16 inputs.

Tolerances: outputs are bf16 (rel 2^-8 per element) of fp32-accumulated sums, so rel-L2 <= 5e-3 against the fp32
reference and max-abs <= a few bf16 ulps of the output scale; wgrad outputs are fp32 (rel-L2 <= 1e-4... bf16 inputs
are exact, only the accumulation order differs).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    import b200moe
    b200moe.load_library()
    return b200moe.ops


def _layout(E, rows_per_expert, pad=128):
    """Hand-built plan: expert e owns rows_per_expert[e] rows, segments padded to `pad` (128, or 256 = CTA-pair layout)."""
    offs = [0]
    for n in rows_per_expert:
        offs.append(offs[-1] + (n + pad - 1) // pad * pad)
    R = offs[-1]
    rows_cap = R + 256                                       # spare tiles beyond R must be skipped
    meta = torch.zeros(8, dtype=torch.int32)
    meta[0], meta[1] = R, R // 128
    tile_expert = torch.full((rows_cap // 128,), -1, dtype=torch.int32)
    valid = torch.zeros(rows_cap, dtype=torch.bool)
    row_expert = torch.zeros(rows_cap, dtype=torch.long)
    for e, n in enumerate(rows_per_expert):
        tile_expert[offs[e] // 128: offs[e + 1] // 128] = e
        valid[offs[e]: offs[e] + n] = True
        row_expert[offs[e]: offs[e + 1]] = e
    return torch.tensor(offs, dtype=torch.int32), meta, tile_expert, valid, row_expert, R, rows_cap


def _rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


CASES = [
    (8, 256, 1024, [800, 790, 1, 0, 128, 129, 300, 555], 128),     # config A shape, ragged incl. an empty expert
    (4, 128, 256, [5, 200, 0, 130], 128),                          # smallest tile-able shape
    (16, 512, 2048, [400] * 16, 128),                              # config C shape
    (2, 1024, 4096, [3000, 2500], 128),                            # config D shape, BN=256 path
    (4, 512, 1024, [6000, 5000, 0, 8100], 256),                    # CTA-pair layout: every GEMM on cta_group::2 tiles
    (2, 1024, 4096, [3000, 2500], 256),                            # CTA pairs for up / dgrad_down only
    (8, 1024, 2048, [300, 200, 0, 129, 520, 64, 1, 700], 128),     # weight-gradient GEMMs on CTA pairs (ragged K, empty expert)
    (8, 256, 1024, [3000, 2900, 0, 3100, 2500, 1, 3300, 2800], 256),  # A-stationary pairs at d_model = 256 (4 k-blocks resident)
    (16, 512, 2048, [1500] * 15 + [700], 256),                     # A-stationary pairs at d_model = 512, run of 2 N tiles
]


@pytest.mark.parametrize("E,d,F,rows,pad", CASES)
def test_ffn_forward_stages(ops, E, d, F, rows, pad):
    g = torch.Generator().manual_seed(E * 1000 + d)
    offs, meta, te, valid, rexp, R, rows_cap = _layout(E, rows, pad)
    xp = torch.randn(rows_cap, d, generator=g).to(torch.bfloat16)
    xp[~valid] = 0
    w1 = (torch.randn(E, F, d, generator=g) * d ** -0.5).to(torch.bfloat16)
    b1 = torch.randn(E, F, generator=g) * 0.1
    w2 = (torch.randn(E, d, F, generator=g) * F ** -0.5).to(torch.bfloat16)
    b2 = torch.randn(E, d, generator=g) * 0.1
    dev = "cuda"
    pre, h = ops.ffn_up(xp.to(dev), w1.to(dev), b1.to(dev), meta.to(dev), te.to(dev), row_pad=pad)
    yo = ops.ffn_down(h, w2.to(dev), b2.to(dev), meta.to(dev), te.to(dev), row_pad=pad)
    torch.cuda.synchronize()
    pre, h, yo = pre.cpu()[:R], h.cpu()[:R], yo.cpu()[:R]
    for e in range(E):
        sl = slice(int(offs[e]), int(offs[e + 1]))
        if sl.stop == sl.start:
            continue
        pre_ref = xp[sl].float() @ w1[e].float().t() + b1[e]
        assert _rel(pre[sl], pre_ref) < 5e-3, f"pre expert {e}"
        h_ref = torch.nn.functional.gelu(pre_ref)
        assert _rel(h[sl], h_ref) < 6e-3, f"h expert {e}"
        yo_ref = h[sl].float() @ w2[e].float().t() + b2[e]          # from the kernel's own bf16 h
        assert _rel(yo[sl], yo_ref) < 5e-3, f"yo expert {e}"
        assert float((yo[sl].float() - yo_ref).abs().max()) < 0.05 * float(yo_ref.abs().max())


@pytest.mark.parametrize("E,d,F,rows,pad", CASES)
def test_ffn_backward_stages(ops, E, d, F, rows, pad):
    g = torch.Generator().manual_seed(E * 1000 + d + 1)
    offs, meta, te, valid, rexp, R, rows_cap = _layout(E, rows, pad)
    dev = "cuda"
    dyp = torch.randn(rows_cap, d, generator=g).to(torch.bfloat16)
    dyp[~valid] = 0
    pre = torch.randn(rows_cap, F, generator=g).to(torch.bfloat16)
    h = torch.randn(rows_cap, F, generator=g).to(torch.bfloat16)
    xp = torch.randn(rows_cap, d, generator=g).to(torch.bfloat16)
    xp[~valid] = 0
    w1 = (torch.randn(E, F, d, generator=g) * d ** -0.5).to(torch.bfloat16)
    w2 = (torch.randn(E, d, F, generator=g) * F ** -0.5).to(torch.bfloat16)
    md, ted, od = meta.to(dev), te.to(dev), offs.to(dev)

    dhpre = ops.ffn_dgrad_down(dyp.to(dev), w2.to(dev), pre.to(dev), md, ted, row_pad=pad)
    dxp = ops.ffn_dgrad_up(dhpre, w1.to(dev), md, ted, row_pad=pad)
    dw2 = ops.ffn_wgrad(dyp.to(dev), h.to(dev), od, E)
    dw1 = ops.ffn_wgrad(dhpre, xp.to(dev), od, E)
    dw2m, dw1m = ops.ffn_wgrad2(dyp.to(dev), h.to(dev), dhpre, xp.to(dev), od, E)    # both in one launch: bit-identical
    assert torch.equal(dw2m, dw2) and torch.equal(dw1m, dw1)
    db2 = ops.colsum(dyp.to(dev), od, E)
    db1 = ops.colsum(dhpre, od, E)
    torch.cuda.synchronize()
    dhpre_c, dxp_c = dhpre.cpu(), dxp.cpu()
    for e in range(E):
        sl = slice(int(offs[e]), int(offs[e + 1]))
        if sl.stop == sl.start:
            assert float(dw2[e].abs().max()) == 0.0 and float(dw1[e].abs().max()) == 0.0
            assert float(db2[e].abs().max()) == 0.0
            continue
        pf = pre[sl].float().requires_grad_(True)
        torch.nn.functional.gelu(pf).backward(dyp[sl].float() @ w2[e].float())
        assert _rel(dhpre_c[sl], pf.grad) < 6e-3, f"dhpre expert {e}"
        dxp_ref = dhpre_c[sl].float() @ w1[e].float()
        assert _rel(dxp_c[sl], dxp_ref) < 5e-3, f"dxp expert {e}"
        dw2_ref = dyp[sl].float().t() @ h[sl].float()
        assert _rel(dw2[e].cpu(), dw2_ref) < 1e-4, f"dw2 expert {e}"
        dw1_ref = dhpre_c[sl].float().t() @ xp[sl].float()
        assert _rel(dw1[e].cpu(), dw1_ref) < 1e-4, f"dw1 expert {e}"
        assert torch.allclose(db2[e].cpu(), dyp[sl].float().sum(0), atol=1e-2, rtol=1e-3)
        assert torch.allclose(db1[e].cpu(), dhpre_c[sl].float().sum(0), atol=1e-2, rtol=1e-3)


@pytest.mark.parametrize("E,d,F,rows,pad", CASES)
def test_gathered_operands_equal_permuted_copy(ops, E, d, F, rows, pad):
    """The dispatch gather fused into the TMA operand loads (tile::gather4 through token_of_row; pads = -1 -> zeros):
    the up-projection and the weight gradient of w1 must be bit-identical to the same GEMMs on the permuted copy."""
    g = torch.Generator().manual_seed(E * 77 + d)
    offs, meta, te, valid, rexp, R, rows_cap = _layout(E, rows, pad)
    dev = "cuda"
    T = max(int(valid.sum()) // 2, 8)                                # every token is used about twice (top-2)
    x = torch.randn(T, d, generator=g).to(torch.bfloat16)
    tok = torch.full((rows_cap,), -1, dtype=torch.int32)
    tok[valid] = torch.randint(0, T, (int(valid.sum()),), generator=g, dtype=torch.int32)
    xp = torch.zeros(rows_cap, d, dtype=torch.bfloat16)
    xp[valid] = x[tok[valid].long()]
    w1 = (torch.randn(E, F, d, generator=g) * d ** -0.5).to(torch.bfloat16).to(dev)
    b1 = (torch.randn(E, F, generator=g) * 0.1).to(dev)
    md, ted, od = meta.to(dev), te.to(dev), offs.to(dev)
    # NaN poison in the freed blocks the outputs will reuse: rows the kernel must not leave unwritten show up
    junk = torch.full((rows_cap * F,), float("nan"), dtype=torch.bfloat16, device=dev)
    del junk
    pre0, h0 = ops.ffn_up(xp.to(dev), w1, b1, md, ted, row_pad=pad)
    clear = torch.full((4,), 7, dtype=torch.int32, device=dev)
    pre1, h1 = ops.ffn_up_gather(x.to(dev), tok.to(dev), w1, b1, md, ted, rows_cap, row_pad=pad, clear=clear)
    torch.cuda.synchronize()
    assert torch.equal(pre0[:R], pre1[:R]) and torch.equal(h0[:R], h1[:R])
    assert clear.tolist() == [0, 7, 7, 7]
    dh = torch.randn(rows_cap, F, generator=g).to(torch.bfloat16)
    dh[~valid] = 0
    dw_ref = ops.ffn_wgrad(dh.to(dev), xp.to(dev), od, E)
    dw = ops.ffn_wgrad_gather(dh.to(dev), x.to(dev), tok.to(dev), od, E)
    torch.cuda.synchronize()
    assert torch.equal(dw, dw_ref)


def test_wgrad_accumulate(ops):
    E, M, N = 2, 128, 256
    offs, meta, te, valid, rexp, R, rows_cap = _layout(E, [100, 260])
    g = torch.Generator().manual_seed(9)
    a = torch.randn(rows_cap, M, generator=g).to(torch.bfloat16)
    b = torch.randn(rows_cap, N, generator=g).to(torch.bfloat16)
    a[~valid] = 0
    od = offs.cuda()
    out = ops.ffn_wgrad(a.cuda(), b.cuda(), od, E)
    out2 = ops.ffn_wgrad(a.cuda(), b.cuda(), od, E, out=out.clone(), accumulate=True)
    assert torch.allclose(out2, 2 * out, rtol=1e-6, atol=1e-6)


def test_bad_shapes_fail_loudly(ops):
    offs, meta, te, valid, rexp, R, rows_cap = _layout(2, [10, 10])
    xp = torch.zeros(rows_cap, 96, dtype=torch.bfloat16, device="cuda")
    w1 = torch.zeros(2, 256, 96, dtype=torch.bfloat16, device="cuda")
    b1 = torch.zeros(2, 256, device="cuda")
    with pytest.raises(RuntimeError, match="multiple of 128"):
        ops.ffn_up(xp, w1, b1, meta.cuda(), te.cuda())


def test_gelu_saturates_for_large_preactivations(ops):
    """|pre| far beyond 8 (where the tanh-form fit of Phi is clamped): h -> pre or 0, GELU' -> 1 or 0, nothing NaN."""
    E, d, F, rows, pad = 2, 128, 256, [256, 128], 128
    g = torch.Generator().manual_seed(77)
    offs, meta, te, valid, rexp, R, rows_cap = _layout(E, rows, pad)
    xp = (torch.randn(rows_cap, d, generator=g) * 6.0).to(torch.bfloat16)
    xp[~valid] = 0
    w1 = (torch.randn(E, F, d, generator=g) * d ** -0.5 * 4.0).to(torch.bfloat16)      # pre ~ N(0, 24^2)
    b1 = torch.zeros(E, F)
    dev = "cuda"
    pre, h = ops.ffn_up(xp.to(dev), w1.to(dev), b1.to(dev), meta.to(dev), te.to(dev), row_pad=pad)
    pre_c, h_c = pre.cpu()[:R].float(), h.cpu()[:R].float()
    assert float(pre_c.abs().max()) > 40 and torch.isfinite(h_c).all()
    h_ref = torch.nn.functional.gelu(pre_c)
    assert _rel(h_c, h_ref) < 6e-3
    big = pre_c.abs() > 9
    assert torch.equal(h_c[big & (pre_c > 0)], pre_c[big & (pre_c > 0)]) and float(h_c[big & (pre_c < 0)].abs().max()) == 0.0
    w2 = (torch.randn(E, d, F, generator=g) * F ** -0.5).to(torch.bfloat16)
    dyp = torch.randn(rows_cap, d, generator=g).to(torch.bfloat16)
    dyp[~valid] = 0
    dhpre = ops.ffn_dgrad_down(dyp.to(dev), w2.to(dev), pre, meta.to(dev), te.to(dev), row_pad=pad).cpu()[:R].float()
    assert torch.isfinite(dhpre).all()
    for e in range(E):
        sl = slice(int(offs[e]), int(offs[e + 1]))
        pf = pre_c[sl].clone().requires_grad_(True)
        torch.nn.functional.gelu(pf).backward(dyp[sl].float() @ w2[e].float())
        assert _rel(dhpre[sl], pf.grad) < 6e-3
