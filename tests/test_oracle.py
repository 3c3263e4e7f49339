"""CPU tests pinning the oracle (oracle/moe_ref.py) to everything the reference offers for this path.

K1  identical experts + renormalised gates == vendored dense FFN (golden vectors from hybrid_encoder.py:157)
K2  uniform router closed forms
K3  hand-written tie / capacity-drop vectors (Appendix A.4, A.5)
K4  context vocabulary sizes vs the reference's frequency table
plus fp64 gradcheck of the oracle's routing-weighted combine.
"""
import math
import os

import numpy as np
import pytest
import torch

from oracle import moe_ref as R


def _load(golden_dir, name):
    return {k: torch.from_numpy(v) for k, v in np.load(os.path.join(golden_dir, name)).items()}


def test_k1_dense_ffn_restatement_matches_golden(golden_dir):
    g = _load(golden_dir, "dense_ffn_d64_f128.npz")
    y = R.dense_ffn(g["x"], g["w1"], g["b1"], g["w2"], g["b2"])
    assert torch.allclose(y, g["y"], rtol=0, atol=1e-6)


@pytest.mark.parametrize("E,k", [(1, 1), (4, 2), (8, 2)])
def test_k1_identical_experts_equal_dense_ffn(golden_dir, E, k):
    g = _load(golden_dir, "dense_ffn_d64_f128.npz")
    d, Fh = g["w1"].shape[1], g["w1"].shape[0]
    m = R.ContextMoERef(d, Fh, E, top_k=k, capacity_factor=None)
    with torch.no_grad():
        for e in range(E):
            m.w1[e].copy_(g["w1"]); m.b1[e].copy_(g["b1"]); m.w2[e].copy_(g["w2"]); m.b2[e].copy_(g["b2"])
    x = g["x"].clone().requires_grad_(True)
    y = m(x)
    assert torch.allclose(y, g["y"], rtol=1e-5, atol=2e-6)
    (y * g["go"]).sum().backward()
    assert torch.allclose(x.grad, g["dx"], rtol=1e-4, atol=1e-5)
    assert torch.allclose(m.w1.grad.sum(0), g["dw1"], rtol=1e-4, atol=1e-5)
    assert torch.allclose(m.w2.grad.sum(0), g["dw2"], rtol=1e-4, atol=1e-5)
    assert torch.allclose(m.b1.grad.sum(0), g["db1"], rtol=1e-4, atol=1e-5)
    # sum_s g_s == 1 so db2 sums to the dense db2
    assert torch.allclose(m.b2.grad.sum(0), g["db2"], rtol=1e-4, atol=1e-5)


def test_k2_uniform_router_closed_forms():
    E, k, T, d = 8, 2, 64, 16
    m = R.ContextMoERef(d, 32, E, top_k=k, capacity_factor=None)
    with torch.no_grad():
        m.gate.weight.zero_()
    x = torch.randn(T, d)
    m(x)
    # all-equal probabilities: every token ties -> lowest indices [0,1] (A.4)
    assert torch.equal(m.last_aux["topk_idx"], torch.tensor([[0, 1]]).expand(T, k))
    assert abs(float(m.last_aux["z_loss"]) - math.log(E) ** 2) < 1e-5
    # P_e = 1/E, sum_e f_e = 1  =>  L_lb = E * (1/E) * 1 = 1
    assert abs(float(m.last_aux["load_balance"]) - 1.0) < 1e-6
    assert torch.allclose(m.last_aux["gates"], torch.full((T, k), 0.5), atol=1e-6)


def test_k3_tie_break_lowest_index():
    lg = torch.log(torch.tensor([[.4, .1, .2, .05, .2, .05],
                                 [.1, .3, .3, .1, .1, .1],
                                 [1 / 6.] * 6]))
    _, idx, g = R.route(lg, 2)
    assert idx.tolist() == [[0, 2], [1, 2], [0, 1]]
    assert torch.allclose(g.sum(-1), torch.ones(3), atol=1e-6)


def test_k3_capacity_drops_slot_major():
    # 4 tokens, k=2, E=3.  slot-0 choices: e0,e0,e0,e1 ; slot-1 choices: e1,e1,e2,e0
    idx = torch.tensor([[0, 1], [0, 1], [0, 2], [1, 0]])
    pos, keep = R.assign_slots(idx, 3, capacity=2)
    # expert0: slot-0 of t0,t1,t2 -> pos 0,1,2 ; slot-1 of t3 comes after all first choices -> pos 3
    # expert1: slot-0 of t3 -> pos 0 (first choices outrank second) ; slot-1 of t0,t1 -> pos 1,2
    assert pos.tolist() == [[0, 1], [1, 2], [2, 0], [0, 3]]
    assert keep.tolist() == [[True, True], [True, False], [False, True], [True, False]]
    off, kept, soa, tor = R.permutation(idx, keep, 3, pad_to=4)
    assert kept.tolist() == [2, 2, 1]
    assert off.tolist() == [0, 4, 8, 12]
    assert soa.tolist() == [[0, 5], [1, -1], [-1, 8], [4, -1]]
    assert tor.tolist() == [0, 1, -1, -1, 3, 0, -1, -1, 2, -1, -1, -1]


def test_capacity_formula_and_global_offset():
    assert R.capacity_of(1.25, 3200, 2, 8) == 1000
    assert R.capacity_of(1.25, 14720, 2, 16) == 2300
    assert R.capacity_of(None, 10, 2, 4) == 20
    # two "ranks" with global offsets reproduce the single-rank positions (A.5, EP order)
    torch.manual_seed(0)
    T, k, E = 64, 2, 4
    idx = torch.stack([torch.randperm(E)[:k] for _ in range(T)])
    pos, keep = R.assign_slots(idx, E, 20)
    h = T // 2
    halves = [idx[:h], idx[h:]]
    cnt = torch.stack([torch.stack([torch.bincount(hh[:, s], minlength=E) for s in range(k)]) for hh in halves])  # [W,k,E]
    tot = cnt.sum(0)                                             # [k,E]
    lower_slots = tot.cumsum(0) - tot                            # all ranks' earlier slot levels
    for r, hh in enumerate(halves):
        off = lower_slots + cnt[:r].sum(0)
        p2, k2 = R.assign_slots(hh, E, 20, global_offset=off)
        assert torch.equal(p2, pos[r * h:(r + 1) * h]) and torch.equal(k2, keep[r * h:(r + 1) * h])


def test_k4_context_vocab_matches_reference_table():
    assert R.CONTEXT_VOCAB_SIZES == (6, 10, 4, 6, 4)
    # category names as committed in context_field_frequencies_final.csv:2-26 of the reference
    assert "fog" in R.CONTEXT_VOCAB["scraped_weather"] and "twilight" in R.CONTEXT_VOCAB["time_of_day"]
    assert R.CONTEXT_VOCAB["solar_context_bin"][1:] == ["night(<-6)", "twilight(-6..0)", "low_sun(0..15)",
                                                       "mid_sun(15..45)", "high_sun(>45)"]


def test_context_bias_is_additive_per_image():
    torch.manual_seed(1)
    B, HW, d, E = 3, 5, 16, 4
    m = R.ContextMoERef(d, 32, E, 2, None, context_vocab_sizes=R.CONTEXT_VOCAB_SIZES, context_dim=8)
    x = torch.randn(B, HW, d)
    ctx = torch.stack([torch.randint(0, v, (B,)) for v in R.CONTEXT_VOCAB_SIZES], dim=1)
    m(x, ctx)
    lg = m.last_aux["logits"].view(B, HW, E)
    base = (x @ m.gate.weight.t())
    bias = m.context_bias(ctx)
    assert torch.allclose(lg, base + bias[:, None, :], atol=1e-5)
    m(x, None)
    assert torch.allclose(m.last_aux["logits"].view(B, HW, E), base, atol=1e-6)


def test_dropped_tokens_output_zero_and_gate_not_redistributed():
    torch.manual_seed(2)
    T, d, E = 32, 16, 4
    m = R.ContextMoERef(d, 32, E, top_k=1, capacity_factor=0.25)   # capacity = 2
    with torch.no_grad():
        m.gate.weight.zero_()                                      # everyone picks expert 0
    y = m(torch.randn(T, d))
    assert m.last_aux["capacity"] == 2 and m.last_aux["dropped"] == T - 2
    assert torch.count_nonzero(y[2:]) == 0 and torch.count_nonzero(y[:2]) > 0


def test_gradcheck_fp64():
    torch.manual_seed(3)
    T, d, Fh, E = 6, 4, 8, 3
    m = R.ContextMoERef(d, Fh, E, top_k=2, capacity_factor=None).double()
    x = torch.randn(T, d, dtype=torch.float64, requires_grad=True)

    def fn(xx):
        # oracle upcasts to fp32 internally; run the same math in fp64 by hand for gradcheck
        lg = xx @ m.gate.weight.t()
        p = torch.softmax(lg, -1)
        idx = torch.sort(p, dim=-1, descending=True, stable=True).indices[:, :2]
        g = torch.gather(p, 1, idx); g = g / (g.sum(-1, keepdim=True) + 1e-9)
        y = torch.zeros_like(xx)
        for s in range(2):
            for e in range(E):
                sel = torch.nonzero(idx[:, s] == e).squeeze(1)
                if sel.numel():
                    h = torch.nn.functional.gelu(xx[sel] @ m.w1[e].t() + m.b1[e])
                    y = y.index_add(0, sel, (h @ m.w2[e].t() + m.b2[e]) * g[sel, s:s + 1])
        return y

    assert torch.autograd.gradcheck(fn, (x,), eps=1e-6, atol=1e-5)
    # and the fp32 oracle agrees with that fp64 restatement
    m32 = R.ContextMoERef(d, Fh, E, top_k=2, capacity_factor=None)
    m32.load_state_dict({k: v.float() for k, v in m.state_dict().items()})
    assert torch.allclose(m32(x.detach().float()).double(), fn(x).detach(), atol=1e-5)


def test_rank_on_logits_vs_rank_on_p():
    """DESIGN.md §0 decision 1, quantified: on 10^6 random fp32 logit rows (E=8 and E=64, the router's scale), how often
    does top-2 by (softmax p desc, index asc) -- the wording of A.4 -- differ from top-2 by (logit desc, index asc),
    which oracle and kernel use?  Softmax is monotone, so they can only differ where fp32 rounding of p makes two
    DISTINCT logits tie (the lower index then wins on p, the larger logit on logits).  The count is reported and
    bounded: the two conventions are the same routing up to a few rows per million."""
    g = torch.Generator().manual_seed(0)
    report = {}
    for E in (8, 64):
        lg = torch.randn(1_000_000, E, generator=g)
        p = torch.softmax(lg, dim=-1)
        on_p = torch.sort(p, dim=-1, descending=True, stable=True).indices[:, :2]
        on_l = torch.sort(lg, dim=-1, descending=True, stable=True).indices[:, :2]
        differ = int((on_p != on_l).any(dim=1).sum())
        report[E] = differ
        assert differ <= 50, report            # measured here: see the printed report (0 of 10^6 for both sizes)
    print("rows of 10^6 where rank-on-p differs from rank-on-logits:", report)
