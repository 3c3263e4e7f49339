"""GPU parity: router, capacity plan, permute / combine against the oracle (bit-exact for indices and row maps).

Routing parity is defined on identical fp32 logits: the kernel's own logits are fed through the
oracle's route(); the logits themselves are compared with a tolerance.
"""
import pytest
import torch

from oracle import moe_ref as R

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    import b200moe
    b200moe.load_library()
    return b200moe.ops


def _mk(T, d, E, seed=0, zero_gate=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(T, d, generator=g).to(torch.bfloat16)
    wg = torch.randn(E, d, generator=g) * d ** -0.5
    if zero_gate:
        wg.zero_()
    return x, wg


@pytest.mark.parametrize("T,d,E,k", [(3200, 256, 8, 2), (1000, 512, 16, 2), (777, 1024, 64, 2), (130, 128, 4, 1),
                                     (513, 256, 5, 3), (4096, 256, 32, 4), (777, 512, 64, 8), (300, 128, 9, 8), (1, 128, 2, 2)])
def test_router_matches_oracle(ops, T, d, E, k):
    x, wg = _mk(T, d, E, seed=T)
    B = 1 if T % 8 else 8
    cb = torch.randn(B, E, generator=torch.Generator().manual_seed(1)) * 0.3
    r = ops.router_fwd(x.cuda(), wg.cuda(), cb.cuda(), T // B, k, True)
    torch.cuda.synchronize()
    lg_ref = x.float() @ wg.t() + cb.repeat_interleave(T // B, dim=0)
    lg = r.logits.cpu()
    assert torch.allclose(lg, lg_ref, atol=2e-4, rtol=1e-4)
    p, idx, g = R.route(lg, k, True)                       # oracle routing on the kernel's own logits
    assert torch.equal(r.topk_idx.cpu().long(), idx)       # bit-exact expert choice and slot order
    assert torch.allclose(r.topk_gate.cpu(), g, atol=1e-6, rtol=1e-5)
    # per-block counts and statistics
    import b200moe
    tb = b200moe.load_library().b200moe_router_block(T)
    nblk = (T + tb - 1) // tb
    bc = r.block_counts.cpu()
    assert bc.shape[0] == nblk
    for b in (0, nblk - 1):
        sl = idx[b * tb:(b + 1) * tb]
        for s in range(k):
            assert torch.equal(bc[b, s].long(), torch.bincount(sl[:, s], minlength=E))
    ps = r.block_stats.cpu()
    assert torch.allclose(ps[:, :E].sum(0), p.sum(0), rtol=1e-4, atol=1e-3)
    assert torch.allclose(ps[:, E].sum(), torch.logsumexp(lg, -1).pow(2).sum(), rtol=1e-4)


def test_router_all_ties_pick_lowest_index(ops):
    T, d, E, k = 300, 256, 8, 2
    x, wg = _mk(T, d, E, zero_gate=True)
    r = ops.router_fwd(x.cuda(), wg.cuda(), None, T, k, True)
    assert torch.equal(r.topk_idx.cpu(), torch.tensor([[0, 1]], dtype=torch.int32).expand(T, k))
    assert torch.allclose(r.topk_gate.cpu(), torch.full((T, k), 0.5), atol=1e-6)
    # partial tie at the k-th place: context bias lifts experts 2 and 5 equally
    cb = torch.zeros(1, E)
    cb[0, 2] = cb[0, 5] = 1.0
    r = ops.router_fwd(x.cuda(), wg.cuda(), cb.cuda(), T, 3, False)
    assert torch.equal(r.topk_idx.cpu(), torch.tensor([[2, 5, 0]], dtype=torch.int32).expand(T, 3))


@pytest.mark.parametrize("T,E,k,cf", [(3200, 8, 2, None), (3200, 8, 2, 1.25), (1000, 16, 2, 0.5), (777, 64, 2, 1.0),
                                      (130, 4, 1, 0.25), (5000, 8, 3, 1.25), (128, 8, 2, 1.25), (1, 8, 2, None),
                                      (40000, 8, 2, 1.25), (20000, 16, 2, 1.0), (9000, 64, 4, 1.0), (8321, 40, 3, 0.75), (777, 64, 8, 1.0), (8300, 64, 8, 0.5)])
def test_plan_matches_oracle(ops, T, E, k, cf):
    d = 256
    x, wg = _mk(T, d, E, seed=7 + T)
    wg[: max(1, E // 8)] *= 3.0                              # skew the router so that capacity bites
    r = ops.router_fwd(x.cuda(), wg.cuda(), None, T, k, True)
    idx = r.topk_idx.cpu().long()
    cap = R.capacity_of(cf, T, k, E)
    p = ops.plan(r, T, k, E, cap, 1e-2, 1e-3)
    torch.cuda.synchronize()
    pos, keep = R.assign_slots(idx, E, cap)
    off, kept, soa, tor = R.permutation(idx, keep, E, pad_to=128)
    assert torch.equal(p.expert_offsets.cpu().long(), off)
    assert torch.equal(p.kept_per_expert.cpu().long(), kept)
    assert torch.equal(p.tokens_per_expert.cpu().long(), torch.bincount(idx.reshape(-1), minlength=E))
    assert torch.equal(p.row_of_assignment.cpu().long(), soa)             # bit-exact drops and rows
    Rr = int(off[-1])
    meta = p.meta.cpu()
    assert int(meta[0]) == Rr and int(meta[1]) == Rr // 128
    assert int(meta[2]) == int((~keep).sum()) and int(meta[3]) == int(keep.sum())
    assert torch.equal(p.token_of_row.cpu().long()[:Rr], tor)
    te = p.tile_expert.cpu()
    for e in range(E):
        assert (te[int(off[e]) // 128: int(off[e + 1]) // 128] == e).all()
    assert (te[Rr // 128:] == -1).all()
    # gates ride along with their rows
    gr = p.gate_of_row.cpu()[:Rr]
    g = r.topk_gate.cpu()
    assert torch.equal(gr[soa[keep]], g[keep])
    # losses (A.8-A.10)
    lg = r.logits.cpu()
    pr, _, _ = R.route(lg, k, True)
    lb, z, _ = R.aux_losses(lg, pr, idx, E)
    ls = p.losses.cpu()
    assert abs(float(ls[1]) - float(lb)) < 1e-4 and abs(float(ls[2]) - float(z)) < 1e-3 * max(1.0, float(z))
    assert abs(float(ls[0]) - (1e-2 * float(lb) + 1e-3 * float(z))) < 1e-4


def test_plan_global_offset_matches_single_rank(ops):
    """Two 'ranks' with global offsets reproduce the single-rank drops (A.5 under expert parallelism)."""
    T, d, E, k = 2048, 256, 8, 2
    x, wg = _mk(T, d, E, seed=3)
    wg[0] *= 4.0
    full = ops.router_fwd(x.cuda(), wg.cuda(), None, T, k, True)
    idx = full.topk_idx.cpu().long()
    cap = R.capacity_of(1.0, T, k, E)
    pos, keep = R.assign_slots(idx, E, cap)
    h = T // 2
    halves = [idx[:h], idx[h:]]
    cnt = torch.stack([torch.stack([torch.bincount(hh[:, s], minlength=E) for s in range(k)]) for hh in halves])
    tot = cnt.sum(0)
    lower = tot.cumsum(0) - tot
    for rnk in range(2):
        xr = x[rnk * h:(rnk + 1) * h].cuda()
        rr = ops.router_fwd(xr, wg.cuda(), None, h, k, True)
        off = (lower + cnt[:rnk].sum(0)).to(torch.int32).cuda()
        p = ops.plan(rr, h, k, E, cap, 0.0, 0.0, global_offset=off, row_pad=1,
                     rows_cap=ops.rows_upper_bound(h, k, E, -1))
        kept_local = (p.row_of_assignment.cpu() >= 0)
        assert torch.equal(kept_local, keep[rnk * h:(rnk + 1) * h])
        # send layout: unpadded, experts back to back, rows in priority order
        ko = keep[rnk * h:(rnk + 1) * h]
        kept_cnt = torch.stack([(ko & (halves[rnk] == e)).sum() for e in range(E)])
        assert torch.equal(p.kept_per_expert.cpu().long(), kept_cnt)
        assert torch.equal(p.expert_offsets.cpu().long()[1:], kept_cnt.cumsum(0))


def test_permute_combine_bit_exact(ops):
    T, d, E, k = 1500, 256, 8, 2
    x, wg = _mk(T, d, E, seed=11)
    xc = x.cuda()
    r = ops.router_fwd(xc, wg.cuda(), None, T, k, True)
    cap = R.capacity_of(1.0, T, k, E)
    p = ops.plan(r, T, k, E, cap, 0.0, 0.0)
    xp = ops.permute(xc, p.token_of_row, p.meta, p.rows_cap)
    Rr = int(p.meta[0])
    tor = p.token_of_row.cpu().long()[:Rr]
    ref = torch.zeros(Rr, d, dtype=torch.bfloat16)
    ref[tor >= 0] = x[tor[tor >= 0]]
    assert torch.equal(xp.cpu()[:Rr], ref)                                   # bit-exact dispatch, zero pads
    # combine: y[t] = sum_s g * yo[row] in slot order, fp32 accumulate
    yo = torch.randn(p.rows_cap, d, generator=torch.Generator().manual_seed(5)).to(torch.bfloat16)
    y = ops.combine(yo.cuda(), p.row_of_assignment, r.topk_gate).cpu()
    soa = p.row_of_assignment.cpu().long()
    g = r.topk_gate.cpu()
    acc = torch.zeros(T, d)
    for s in range(k):
        m = soa[:, s] >= 0
        acc[m] = torch.addcmul(acc[m], yo[soa[m, s]].float(), g[m, s:s + 1])   # same fma order as the kernel
    assert torch.allclose(y.float(), acc.to(torch.bfloat16).float(), atol=0, rtol=2 ** -7)
    # backward of combine
    dy = torch.randn(T, d, generator=torch.Generator().manual_seed(6)).to(torch.bfloat16)
    dyp, dgr = ops.combine_bwd(dy.cuda(), yo.cuda(), p.token_of_row, p.gate_of_row, p.meta, p.rows_cap)
    gor = p.gate_of_row.cpu()[:Rr]
    ref = torch.zeros(Rr, d)
    ref[tor >= 0] = dy[tor[tor >= 0]].float() * gor[tor >= 0, None]
    assert torch.equal(dyp.cpu()[:Rr], ref.to(torch.bfloat16))
    dref = torch.zeros(Rr)
    dref[tor >= 0] = (dy[tor[tor >= 0]].float() * yo[:Rr][tor >= 0].float()).sum(-1)
    assert torch.allclose(dgr.cpu()[:Rr], dref, atol=1e-3, rtol=1e-4)
    # unpermute backward without the router term
    dxp = torch.randn(p.rows_cap, d, generator=torch.Generator().manual_seed(8)).to(torch.bfloat16)
    dx = ops.unpermute_bwd(dxp.cuda(), p.row_of_assignment, None, wg.cuda(), T, k, d).cpu()
    acc = torch.zeros(T, d)
    for s in range(k):
        m = soa[:, s] >= 0
        acc[m] += dxp[soa[m, s]].float()
    assert torch.equal(dx, acc.to(torch.bfloat16))


def test_errors_are_loud(ops):
    x, wg = _mk(64, 100, 8)                                # d not a multiple of 64
    with pytest.raises(RuntimeError, match="multiple of 64"):
        ops.router_fwd(x.cuda(), wg.cuda(), None, 64, 2, True)
    x, wg = _mk(64, 128, 4)
    with pytest.raises(RuntimeError, match="top_k"):
        ops.router_fwd(x.cuda(), wg.cuda(), None, 64, 5, True)
