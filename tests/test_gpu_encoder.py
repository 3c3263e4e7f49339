"""GPU: encoder-layer glue.  A dense stand-in with the attribute names and forward of
the vendored TransformerEncoderLayer (hybrid_encoder.py:116-161) is upcycled in place; with identical experts,
renormalised gates and no capacity limit the swapped layer must reproduce the dense one (K1 anchor) and train."""
import copy
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu


class TransformerEncoderLayer(nn.Module):           # name matters: upcycle_encoder matches the vendored class name
    def __init__(self, d_model, nhead, dim_feedforward, normalize_before=False):
        super().__init__()
        self.normalize_before = normalize_before
        self.self_attn = nn.MultiheadAttention(d_model, nhead, 0.0, batch_first=True)
        self.linear1, self.linear2 = nn.Linear(d_model, dim_feedforward), nn.Linear(dim_feedforward, d_model)
        self.norm1, self.norm2 = nn.LayerNorm(d_model), nn.LayerNorm(d_model)
        self.dropout, self.dropout1, self.dropout2 = nn.Dropout(0.0), nn.Dropout(0.0), nn.Dropout(0.0)
        self.activation = nn.GELU()

    def forward(self, src, src_mask=None, pos_embed=None):
        residual = src
        if self.normalize_before:
            src = self.norm1(src)
        q = k = src if pos_embed is None else src + pos_embed
        src, _ = self.self_attn(q, k, value=src, attn_mask=src_mask)
        src = residual + self.dropout1(src)
        if not self.normalize_before:
            src = self.norm1(src)
        residual = src
        if self.normalize_before:
            src = self.norm2(src)
        src = self.linear2(self.dropout(self.activation(self.linear1(src))))
        src = residual + self.dropout2(src)
        if not self.normalize_before:
            src = self.norm2(src)
        return src


class Detector(nn.Module):
    def __init__(self, layer):
        super().__init__()
        self.encoder = nn.ModuleList([copy.deepcopy(layer)])      # hybrid_encoder.py:167 deep-copies the layer

    def forward(self, x, targets=None):
        for l_ in self.encoder:
            x = l_(x)
        return x


@pytest.mark.parametrize("pre_norm", [False, True])
def test_upcycled_encoder_layer_matches_dense_and_trains(pre_norm):
    import b200moe
    from b200moe import context as cx
    torch.manual_seed(0)
    B, HW, d, F = 4, 400, 256, 1024
    dense = Detector(TransformerEncoderLayer(d, 8, F, pre_norm)).cuda()
    moe = copy.deepcopy(dense)
    n = b200moe.upcycle_encoder(moe, num_experts=8, top_k=2, capacity_factor=None,
                                context_vocab_sizes=cx.CONTEXT_VOCAB_SIZES)
    assert n == 1 and isinstance(moe.encoder[0], b200moe.MoEEncoderLayer)
    hook = cx.install_context_hook(moe)
    x = torch.randn(B, HW, d, device="cuda")
    targets = [cx.attach_context({}, {"time_of_day": "night", "scraped_weather": "rain"}) for _ in range(B)]
    with torch.no_grad():
        yd = dense(x)
    xm = x.clone().requires_grad_(True)
    ym = moe(xm, targets=targets)
    assert moe.encoder[0]._context.shape == (B, 5)
    rel = float((ym.float() - yd).norm() / yd.norm())
    assert rel <= 1e-2, rel                                         # bf16 experts vs the fp32 dense FFN
    aux = cx.collect_aux_loss(moe)
    assert aux is not None and aux.requires_grad
    (ym.float().pow(2).mean() + aux).backward()
    layer = moe.encoder[0]
    for p in (layer.moe.w1, layer.moe.w2, layer.moe.gate.weight, layer.moe.ctx_proj.weight, layer.norm2.weight):
        assert p.grad is not None and torch.isfinite(p.grad).all() and p.grad.abs().sum() > 0
    assert xm.grad is not None and torch.isfinite(xm.grad).all()
    tel = cx.router_telemetry(moe)
    assert tel["moe0/dropped"] == 0 and 0.0 < tel["moe0/max_expert_share"] <= 1.0
    hook.remove()
    copy.deepcopy(moe)                                              # EMA / hybrid_encoder.py:167 deep-copy the module


def test_decoder_ffn_site_shape():
    """§8(f) rank 4: the decoder FFN (rtdetrv2_decoder.py:186-189: d=256, ffn=1024, 300 queries per image) is the same
    layer at a smaller token count; per-expert row counts fall below one 128-row tile."""
    import b200moe
    from oracle import moe_ref as R
    torch.manual_seed(1)
    B, Q, d, F, E, k = 8, 300, 256, 1024, 8, 2
    ref = R.ContextMoERef(d, F, E, k, 1.25)
    with torch.no_grad():
        for p in (ref.w1, ref.w2):
            p.copy_(p.to(torch.bfloat16).float())
    m = b200moe.ContextMoE(d, F, E, k, 1.25)
    m.load_state_dict(ref.state_dict())
    m = m.cuda()
    x = torch.randn(B, Q, d).to(torch.bfloat16)
    y = m(x.cuda())
    yr = ref(x.float(), None, logits_value=m.last_aux["logits"].cpu())
    assert torch.equal(m.last_aux["topk_idx"].cpu().long(), ref.last_aux["topk_idx"])
    assert float((y.float().cpu() - yr).norm() / yr.norm()) <= 1e-2


@pytest.mark.parametrize("B,HW,d,F,E,cf", [(8, 400, 256, 1024, 8, 1.25), (3, 231, 512, 1024, 16, 0.5), (2, 100, 128, 256, 4, None),
                                           (2, 77, 1024, 1024, 8, 1.0)])
@pytest.mark.parametrize("combine", ["auto", "tail", "epilogue", "separate"])
def test_fused_post_norm_matches_unfused(B, HW, d, F, E, cf, combine):
    """LayerNorm(residual + MoE(x)) through the fused combine kernel (forward and backward) against the same module
    with torch's add + LayerNorm around the plain forward.  Both run the same expert kernels on the same routing, so
    the only differences are the bf16 rounding of y that the fused path skips and summation order."""
    import b200moe
    torch.manual_seed(B * HW)
    m = b200moe.ContextMoE(d, F, E, 2, cf, context_vocab_sizes=(6, 10, 4, 6, 4)).cuda()
    m.combine = combine          # LayerNorm in the down-projection's tail (both fused modes) or in its own kernel
    norm = nn.LayerNorm(d).cuda()
    with torch.no_grad():
        norm.weight.uniform_(0.5, 1.5)
        norm.bias.uniform_(-0.2, 0.2)
    x = torch.randn(B, HW, d, device="cuda").to(torch.bfloat16)
    res = torch.randn(B, HW, d, device="cuda").to(torch.bfloat16)
    go = torch.randn(B, HW, d, device="cuda")
    ctx = torch.stack([torch.randint(0, v, (B,)) for v in (6, 10, 4, 6, 4)], 1).cuda()

    def run(fused):
        for p in list(m.parameters()) + list(norm.parameters()):
            p.grad = None
        xa, ra = x.clone().requires_grad_(True), res.clone().requires_grad_(True)
        out = m.forward_post_norm(xa, ra, norm, ctx) if fused else norm((ra + m(xa, ctx)).float()).to(torch.bfloat16)
        ((out.float() * go).sum() + m.aux_loss).backward()
        grads = {n: p.grad.clone() for n, p in list(m.named_parameters()) + [("ln." + n_, p_) for n_, p_ in norm.named_parameters()]}
        return out.detach(), xa.grad, ra.grad, grads

    of, dxf, drf, gf = run(True)
    ou, dxu, dru, gu = run(False)

    def rel(a, b):
        return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))
    assert of.dtype == torch.bfloat16 and rel(of, ou) <= 6e-3
    assert rel(dxf, dxu) <= 1e-2 and rel(drf, dru) <= 1e-2
    for n in gu:
        assert rel(gf[n], gu[n]) <= 2e-2, n
    # deterministic
    of2, dxf2, _, gf2 = run(True)
    assert torch.equal(of, of2) and torch.equal(dxf, dxf2) and torch.equal(gf["ln.weight"], gf2["ln.weight"])


def _golden_layer(golden_dir, name):
    g = {k: torch.from_numpy(v) for k, v in np.load(os.path.join(golden_dir, name)).items()}
    sd = {k[3:]: v for k, v in g.items() if k.startswith("sd.")}
    d, F = sd["linear1.weight"].shape[1], sd["linear1.weight"].shape[0]
    dense = TransformerEncoderLayer(d, 8, F, False)
    dense.load_state_dict(sd)                      # the vendored layer's own state_dict keys (hybrid_encoder.py:126-141)
    return g, dense


@pytest.mark.parametrize("fp32_stream", [False, True])
def test_fused_post_norm_matches_vendored_golden(golden_dir, fp32_stream):
    """Reference-derived known answer for §8(f)2: `norm2(src + linear2(gelu(linear1(src))))` computed by the UNMODIFIED
    vendored layer (tests/golden/make_golden.py, hybrid_encoder.py:157-160) against ContextMoE.forward_post_norm with E
    identical experts (renormalised gates, no drops), forward and backward.  fp32_stream: residual and output stay
    fp32, as the encoder's residual stream is under the reference's autocast (LayerNorm outputs fp32)."""
    import b200moe
    g, dense = _golden_layer(golden_dir, "postnorm_ffn_d128_f256.npz")
    layer = b200moe.MoEEncoderLayer.from_dense(dense, num_experts=4, top_k=2, capacity_factor=None).cuda()
    src = g["src"].cuda().requires_grad_(True)
    if fp32_stream:
        out = layer.moe.forward_post_norm(src, src, layer.norm2)
        assert out.dtype == torch.float32
    else:
        out = layer.moe.forward_post_norm(src.to(torch.bfloat16), src.to(torch.bfloat16), layer.norm2)
        assert out.dtype == torch.bfloat16
    (out.float() * g["go"].cuda()).sum().backward()

    def rel(a, b):
        return float((a.float().cpu() - b.float()).norm() / (b.float().norm() + 1e-12))
    assert rel(out, g["out"]) <= 1e-2
    assert rel(src.grad, g["dsrc"]) <= 1e-2
    assert rel(layer.norm2.weight.grad, g["dnorm2_weight"]) <= 1e-2
    assert rel(layer.norm2.bias.grad, g["dnorm2_bias"]) <= 1e-2
    assert rel(layer.moe.w1.grad.sum(0), g["dw1"]) <= 1e-2
    assert rel(layer.moe.w2.grad.sum(0), g["dw2"]) <= 1e-2


def test_upcycled_layer_matches_vendored_encoder_layer_golden(golden_dir):
    """The whole vendored TransformerEncoderLayer.forward (hybrid_encoder.py:143-161: attention with pos_embed, both
    residuals, both LayerNorms) against MoEEncoderLayer.from_dense on the same weights, output and input gradient."""
    import b200moe
    g, dense = _golden_layer(golden_dir, "encoder_layer_d128_f256.npz")
    layer = b200moe.MoEEncoderLayer.from_dense(dense.cuda(), num_experts=4, top_k=2, capacity_factor=None)
    src = g["src"].cuda().requires_grad_(True)
    out = layer(src, pos_embed=g["pos"].cuda())
    (out.float() * g["go"].cuda()).sum().backward()
    rel = float((out.float().cpu() - g["out"]).norm() / g["out"].norm())
    reld = float((src.grad.float().cpu() - g["dsrc"]).norm() / g["dsrc"].norm())
    assert rel <= 1e-2 and reld <= 1e-2, (rel, reld)
