"""GPU parity of the small entry points around the layer: context bias (A.2) forward / backward, one-launch weight
preparation, two-tensor column sums, the router's parameter gradients, and the device-side expert-parallel routing
tables against their host twin (b200moe.ep.route_tables, the function the gloo tests pin on CPU)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    import b200moe
    b200moe.load_library()
    return b200moe.ops


def _call(name, *a):
    from b200moe import _lib
    _lib.call(name, *a)


def test_ctx_bias_matches_torch(ops):
    from b200moe import _lib
    g = torch.Generator().manual_seed(3)
    vocab, cd, B, E = (6, 10, 4, 6, 4), 16, 9, 8
    tabs = [torch.randn(v, cd, generator=g) for v in vocab]
    wc = torch.randn(E, len(vocab) * cd, generator=g) * 0.1
    ctx = torch.stack([torch.randint(0, v, (B,), generator=g) for v in vocab], 1)
    ctx[0, 0] = 0
    dcb = torch.randn(B, E, generator=g)
    # torch reference (fp32)
    tr = [t.clone().requires_grad_(True) for t in tabs]
    wr = wc.clone().requires_grad_(True)
    emb = torch.cat([torch.nn.functional.embedding(ctx[:, i], tr[i]) for i in range(len(vocab))], -1)
    ref = emb @ wr.t()
    ref.backward(dcb)
    # kernel
    dev = "cuda"
    td = [t.to(dev) for t in tabs]
    wd, cdv = wc.to(dev), ctx.to(dev)
    arr = (C.c_void_p * len(td))(*[t.data_ptr() for t in td])
    voc = (C.c_int32 * len(td))(*vocab)
    out = torch.empty(B, E, device=dev)
    st = _lib.stream_ptr()
    _call("b200moe_ctx_bias_fwd", cdv.data_ptr(), C.addressof(arr), C.addressof(voc), len(td), cd, wd.data_ptr(), B, E,
          out.data_ptr(), st)
    assert torch.allclose(out.cpu(), ref.detach(), atol=1e-5, rtol=1e-5)
    dt = [torch.full_like(t, float("nan")) for t in td]
    darr = (C.c_void_p * len(td))(*[t.data_ptr() for t in dt])
    dwc = torch.full_like(wd, float("nan"))
    dd = dcb.to(dev)
    _call("b200moe_ctx_bias_bwd", cdv.data_ptr(), dd.data_ptr(), C.addressof(arr), C.addressof(voc), len(td), cd,
          wd.data_ptr(), B, E, C.addressof(darr), dwc.data_ptr(), st)
    assert torch.allclose(dwc.cpu(), wr.grad, atol=1e-5, rtol=1e-5)
    for a, b in zip(dt, tr):
        assert torch.allclose(a.cpu(), b.grad, atol=1e-5, rtol=1e-5)     # rows of unused ids are written as zeros


def test_prepare_weights_matches_casts(ops):
    from b200moe import _lib
    g = torch.Generator().manual_seed(5)
    E, d, F = 5, 256, 384
    w1, w2 = torch.randn(E, F, d, generator=g).cuda(), torch.randn(E, d, F, generator=g).cuda()
    wg = (torch.randn(E, d, generator=g) * d ** -0.5).cuda()
    w1b, w2b = torch.empty_like(w1, dtype=torch.bfloat16), torch.empty_like(w2, dtype=torch.bfloat16)
    n = _lib.load().b200moe_router_scratch_bytes(d, E) // 2
    ws = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    _call("b200moe_prepare_weights", w1.data_ptr(), w1b.data_ptr(), w1.numel(), w2.data_ptr(), w2b.data_ptr(),
          w2.numel(), wg.data_ptr(), ws.data_ptr(), d, E, _lib.stream_ptr())
    assert torch.equal(w1b, w1.to(torch.bfloat16)) and torch.equal(w2b, w2.to(torch.bfloat16))
    EP = 16
    ws = ws.view(2 * EP, d)
    hi = wg.to(torch.bfloat16)
    assert torch.equal(ws[:E], hi)
    assert torch.equal(ws[EP:EP + E], (wg - hi.float()).to(torch.bfloat16))
    assert ws[E:EP].abs().sum() == 0 and ws[EP + E:].abs().sum() == 0
    # hi + lo reproduces the fp32 gate to ~2^-17 relative
    assert ((ws[:E].float() + ws[EP:EP + E].float() - wg).abs() <= wg.abs() * 2 ** -16 + 1e-30).all()


def test_colsum_two_tensors_one_launch(ops):
    g = torch.Generator().manual_seed(9)
    E, rows = 6, (128, 0, 384, 256, 128, 1024)
    offs = torch.tensor([0] + list(torch.tensor(rows).cumsum(0)), dtype=torch.int32)
    R = int(offs[-1])
    a0 = torch.randn(R, 128, generator=g).to(torch.bfloat16)
    a1 = torch.randn(R, 320, generator=g).to(torch.bfloat16)
    d0, d1 = ops.colsum(a0.cuda(), offs.cuda(), E, a1=a1.cuda())
    for a, dout in ((a0, d0), (a1, d1)):
        ref = torch.stack([a[offs[e]:offs[e + 1]].float().sum(0) for e in range(E)])
        assert torch.allclose(dout.cpu(), ref, atol=1e-3, rtol=1e-4)
    again0, again1 = ops.colsum(a0.cuda(), offs.cuda(), E, a1=a1.cuda())
    assert torch.equal(again0, d0) and torch.equal(again1, d1)            # fixed summation order


@pytest.mark.parametrize("T,d,E,tpi", [(3200, 256, 8, 400), (1000, 512, 16, 1000), (777, 1024, 64, 111), (130, 128, 5, 65),
                                       # from 8192 tokens on the contraction runs on the tensor pipe (bf16 hi/lo split of
                                       # dlogits, row chunks as the experts of the weight-gradient GEMM): ragged last
                                       # chunk, T not a multiple of 64, E below / at the 64-column half of the split
                                       (8192, 256, 8, 400), (9001, 128, 5, 450), (16384, 1024, 64, 1024),
                                       (65533, 512, 16, 923)])
def test_router_param_grads(ops, T, d, E, tpi):
    g = torch.Generator().manual_seed(T)
    dl = torch.randn(T, E, generator=g) * 0.01
    x = torch.randn(T, d, generator=g).to(torch.bfloat16)
    dwg, dctx = ops.router_bwd_params(dl.cuda(), x.cuda(), tpi)
    ref = (dl.double().t() @ x.double()).float()
    assert torch.allclose(dwg.cpu(), ref, atol=1e-4, rtol=1e-4)
    assert float((dwg.cpu() - ref).norm() / ref.norm()) < 2e-5
    B = (T + tpi - 1) // tpi
    refc = torch.stack([dl[b * tpi:(b + 1) * tpi].sum(0) for b in range(B)])
    assert torch.allclose(dctx.cpu(), refc, atol=1e-5, rtol=1e-5)
    dwg2, _ = ops.router_bwd_params(dl.cuda(), x.cuda(), tpi)
    assert torch.equal(dwg, dwg2)                                         # deterministic
    assert torch.allclose(ops.router_bwd_dwg(dl.cuda(), x.cuda()).cpu(), ref, atol=1e-4, rtol=1e-4)
    assert torch.allclose(ops.router_bwd_dctx(dl.cuda(), tpi).cpu(), refc, atol=1e-5, rtol=1e-5)


@pytest.mark.parametrize("W,k,E,cf,chunks", [(2, 2, 8, 1.25, 1), (4, 2, 16, 0.5, 2), (8, 2, 64, 1.25, 1), (8, 3, 8, None, 1),
                                             (1, 1, 4, 1.0, 2)])
def test_ep_route_tables_match_host_twin(ops, W, k, E, cf, chunks):
    from b200moe import _lib
    from b200moe.ep import chunk_layout, route_tables, split_sizes
    g = torch.Generator().manual_seed(W * 100 + E)
    hist = torch.randint(0, 400, (W, k, E), generator=g, dtype=torch.int32)
    hist[:, :, 0] += 900                                       # one hot expert so that the capacity bites
    tokens = hist.sum((1, 2)).to(torch.int32) // k
    allh = torch.cat([hist.reshape(W, -1), tokens[:, None]], 1).contiguous().cuda()
    C_, ELc, _ = chunk_layout(E, W, chunks)
    for rank in range(W):
        goff_ref, cap_ref, kept_ref = route_tables(hist, tokens, rank, cf, k)
        goff = torch.empty(k, E, dtype=torch.int32, device="cuda")
        kept = torch.empty(W * E + 2, dtype=torch.int32, device="cuda")
        recv = torch.empty(C_, W * ELc, dtype=torch.int32, device="cuda")
        _call("b200moe_ep_route_tables", allh.data_ptr(), W, k, E, rank, -1.0 if cf is None else float(cf), ELc,
              goff.data_ptr(), kept.data_ptr(), recv.data_ptr(), _lib.stream_ptr())
        kept = kept.cpu()
        assert int(kept[W * E]) == cap_ref and int(kept[W * E + 1]) == int(tokens.sum())
        assert torch.equal(kept[:W * E].reshape(W, E).long(), kept_ref)
        assert torch.equal(goff.cpu(), goff_ref)
        _, _, recv_ref, _ = split_sizes(kept_ref, rank, W, C_)
        for c in range(C_):
            assert torch.equal(recv[c].cpu(), recv_ref[c])


def test_ep_hist(ops):
    from b200moe import _lib
    g = torch.Generator().manual_seed(1)
    nblk, k, E, T = 37, 2, 16, 4700
    bc = torch.randint(0, 50, (nblk, k, E), generator=g, dtype=torch.int32).cuda()
    out = torch.empty(k * E + 1, dtype=torch.int32, device="cuda")
    _call("b200moe_ep_hist", bc.data_ptr(), nblk, k, E, T, out.data_ptr(), _lib.stream_ptr())
    assert torch.equal(out[:-1].cpu(), bc.sum(0).reshape(-1).cpu().int()) and int(out[-1]) == T


@pytest.mark.parametrize("W,E,pad", [(2, 8, 128), (4, 16, 256), (8, 64, 128), (1, 4, 256)])
def test_ep_p2p_tables_match_host_twin(ops, W, E, pad):
    from b200moe import _lib
    from b200moe.ep_p2p import p2p_layout
    g = torch.Generator().manual_seed(W * 7 + E)
    kept = torch.randint(0, 700, (W, E), generator=g, dtype=torch.int32)
    kept[:, 1] = 0                                               # an expert nobody routes to
    kd = torch.cat([kept.reshape(-1), torch.tensor([0, 0], dtype=torch.int32)]).cuda()
    EL = E // W
    for rank in range(W):
        dest_ref, offs_ref = p2p_layout(kept, rank, pad)
        rows_cap = int(offs_ref[:, -1].max())
        dest = torch.empty(E, dtype=torch.int32, device="cuda")
        offs = torch.empty(EL + 1, dtype=torch.int32, device="cuda")
        meta = torch.empty(8, dtype=torch.int32, device="cuda")
        tile_e = torch.empty(rows_cap // 128, dtype=torch.int32, device="cuda")
        _call("b200moe_ep_p2p_tables", kd.data_ptr(), W, E, rank, rows_cap, pad, dest.data_ptr(), offs.data_ptr(),
              meta.data_ptr(), tile_e.data_ptr(), _lib.stream_ptr())
        assert torch.equal(dest.cpu(), dest_ref) and torch.equal(offs.cpu(), offs_ref[rank])
        m = meta.cpu()
        R = int(offs_ref[rank, -1])
        assert int(m[0]) == R and int(m[1]) == R // 128 and int(m[5]) == 0
        assert int(m[3]) == int(kept[:, rank * EL:(rank + 1) * EL].sum())
        te = tile_e.cpu()
        for t in range(rows_cap // 128):
            want = -1
            if t * 128 < R:
                want = int((offs_ref[rank, 1:] <= t * 128).sum())
                while want + 1 < EL and offs_ref[rank, want + 1] <= t * 128:
                    want += 1
            assert int(te[t]) == want
        # too small a buffer is flagged, not overrun
        if rows_cap > 128:
            _call("b200moe_ep_p2p_tables", kd.data_ptr(), W, E, rank, rows_cap - 128, pad, dest.data_ptr(), offs.data_ptr(),
                  meta.data_ptr(), tile_e.data_ptr(), _lib.stream_ptr())
            assert int(meta.cpu()[5]) == 1
