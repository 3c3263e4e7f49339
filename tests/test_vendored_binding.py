"""CPU, dev box only (skipped where /root/reference is absent, i.e. on the GPU box): the host glue against the REAL
vendored classes — registry (`@register()` / `create()` from a YAML-style dict, workspace.py:15-53, 92-179),
HybridEncoder (hybrid_encoder.py:183-330), RTDETR (rtdetr.py:19-44), CocoDetection's target dict
(coco_dataset.py:43-62) — loaded read-only through the bypass loader of SURVEY.md Appendix B.  No compute runs here
(the layer has no CPU path): structure, deep copy (ema.py:31), state_dict round trips, the context hook and the
auxiliary-loss plumbing are what is checked."""
import copy
import importlib.util
import os
import sys
import types

import pytest
import torch
import torch.nn as nn

REF = "/root/reference/third_party/rtdetr/rtdetrv2_pytorch/src"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="the vendored reference is not on this box")


@pytest.fixture(scope="module")
def vendored():
    pkg = "_rtv2bind"
    for name in (pkg, f"{pkg}.core", f"{pkg}.zoo", f"{pkg}.zoo.rtdetr"):
        m = types.ModuleType(name)
        m.__path__ = []
        sys.modules[name] = m

    def load(modname, path):
        spec = importlib.util.spec_from_file_location(modname, path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[modname] = mod
        spec.loader.exec_module(mod)
        return mod

    ws = load(f"{pkg}.core.workspace", f"{REF}/core/workspace.py")
    sys.modules[f"{pkg}.core"].register = ws.register
    load(f"{pkg}.zoo.rtdetr.utils", f"{REF}/zoo/rtdetr/utils.py")
    he = load(f"{pkg}.zoo.rtdetr.hybrid_encoder", f"{REF}/zoo/rtdetr/hybrid_encoder.py")
    rt = load(f"{pkg}.zoo.rtdetr.rtdetr", f"{REF}/zoo/rtdetr/rtdetr.py")
    return ws, he, rt


def test_registry_builds_moe_encoder_from_yaml_style_config(vendored):
    import b200moe
    ws, he, _ = vendored
    classes = b200moe.registry.register_all(ws.register, he.HybridEncoder)
    assert {"ContextMoE", "MoEEncoderLayer", "MoEHybridEncoder"} <= set(classes)
    cfg = ws.GLOBAL_CONFIG
    # the schema extraction saw every constructor argument of the layer classes (no **kwargs lost)
    assert {"num_experts", "top_k", "capacity_factor", "context_vocab_sizes"} <= set(cfg["MoEEncoderLayer"]["_kwargs"])
    assert {"num_experts", "top_k", "capacity_factor", "aux_loss_weight"} <= set(cfg["ContextMoE"]["_kwargs"])
    assert set(cfg["HybridEncoder"]["_kwargs"]) <= set(cfg["MoEHybridEncoder"]["_kwargs"])
    # what YAMLConfig does with `MoEHybridEncoder: {...}` (yaml_config.py merges the dict into the schema)
    cfg["MoEHybridEncoder"].update({"in_channels": [128, 256, 512], "hidden_dim": 256, "expansion": 0.5,
                                    "num_experts": 4, "top_k": 2, "capacity_factor": 1.5})
    enc = ws.create("MoEHybridEncoder", cfg)
    assert isinstance(enc, he.HybridEncoder) and enc.num_moe_layers == 1
    layer = enc.encoder[0].layers[0]
    assert isinstance(layer, b200moe.MoEEncoderLayer) and not hasattr(layer, "linear1")
    moe = layer.moe
    assert (moe.num_experts, moe.top_k, moe.capacity_factor, moe.activation) == (4, 2, 1.5, "gelu")
    assert moe.context_vocab_sizes == b200moe.context.CONTEXT_VOCAB_SIZES
    # every expert starts as the dense FFN the vendored constructor initialised (K1 anchor): identical slabs
    assert torch.equal(moe.w1[0], moe.w1[3]) and torch.equal(moe.b2[0], moe.b2[2])
    # a single layer / the bare MoE through create() as well
    cfg["MoEEncoderLayer"].update({"d_model": 256, "nhead": 8, "dim_feedforward": 1024, "num_experts": 8})
    assert ws.create("MoEEncoderLayer", cfg).moe.num_experts == 8
    cfg["ContextMoE"].update({"d_model": 256, "dim_feedforward": 512, "num_experts": 16, "activation": "relu"})
    assert ws.create("ContextMoE", cfg).activation == "relu"


def test_upcycled_vendored_encoder_invariants(vendored):
    """upcycle_encoder + install_context_hook on a real HybridEncoder inside a real RTDETR: module structure, deepcopy
    (ModelEMA, ema.py:31), plain state_dict round trip (_solver.py:102-116), parameter names clear of the optimizer's
    norm regex (include/optimizer.yml:18-22), context reaching the layer, the criterion wrapper."""
    import re
    import b200moe
    from b200moe import context as cx
    _, he, rt = vendored
    enc = he.HybridEncoder(in_channels=[128, 256, 512], hidden_dim=256, expansion=0.5)
    dense_sd = copy.deepcopy(enc.state_dict())
    n = b200moe.upcycle_encoder(enc, num_experts=8, top_k=2, capacity_factor=1.25,
                                context_vocab_sizes=cx.CONTEXT_VOCAB_SIZES)
    assert n == 1
    layer = enc.encoder[0].layers[0]
    assert isinstance(layer, b200moe.MoEEncoderLayer) and layer.moe.activation == "gelu"
    # attention and norms are the dense layer's own tensors; the FFN went into the experts
    for k_ in ("self_attn.in_proj_weight", "norm1.weight", "norm2.bias"):
        assert torch.equal(enc.state_dict()[f"encoder.0.layers.0.{k_}"], dense_sd[f"encoder.0.layers.0.{k_}"])
    assert torch.equal(layer.moe.w1[5], dense_sd["encoder.0.layers.0.linear1.weight"])
    assert torch.equal(layer.moe.b2[0], dense_sd["encoder.0.layers.0.linear2.bias"])

    class Stub(nn.Module):
        def forward(self, x, targets=None):
            return x

    det = rt.RTDETR(Stub(), enc, Stub())
    twin = copy.deepcopy(det)                                           # ModelEMA
    assert set(twin.state_dict()) == set(det.state_dict())
    twin.load_state_dict(det.state_dict())                              # resume / EMA update iterate plain tensors
    assert all(isinstance(v, torch.Tensor) for v in det.state_dict().values())
    names = [n_ for n_, _ in det.named_parameters()]
    moe_names = [n_ for n_ in names if ".moe." in n_]
    assert {"encoder.encoder.0.layers.0.moe.gate.weight", "encoder.encoder.0.layers.0.moe.w1"} <= set(moe_names)
    # the two regex parameter groups of configs/rtdetrv2/include/optimizer.yml:14-19 must not capture MoE parameters
    for pattern in (r"^(?=.*backbone)(?!.*norm).*$", r"^(?=.*(?:encoder|decoder))(?=.*(?:norm|bn)).*$"):
        assert not any(re.match(pattern, n_) for n_ in moe_names), pattern

    hook = cx.install_context_hook(det)
    rec = {"solar_context_bin": "night(<-6)", "scraped_weather": "fog", "time_of_day": "night",
           "road_type": "city", "road_condition": "wet"}
    targets = [cx.attach_context({"boxes": torch.zeros(0, 4)}, rec), cx.attach_context({}, {})]
    seen = {}
    orig = layer.forward
    layer.forward = lambda src, src_mask=None, pos_embed=None: seen.setdefault("ctx", layer._context) is None or src
    try:
        enc_in = [torch.zeros(2, c, s, s) for c, s in ((128, 8), (256, 4), (512, 2))]
        det(enc_in, targets=targets)                                    # RTDETR.forward (rtdetr.py:32-37), real encoder
    finally:
        layer.forward = orig
        hook.remove()
    assert seen["ctx"].tolist() == [cx.context_ids(rec), [0] * 5]

    class Crit(nn.Module):
        def forward(self, outputs, targets, **kw):
            return {"loss_vfl": torch.ones(())}

    wrapped = b200moe.registry.MoECriterion(Crit(), det)
    assert set(wrapped(None, None)) == {"loss_vfl"}                     # no forward yet: nothing to add
    layer.moe.aux_loss = torch.tensor(0.25)                             # (a CUDA scalar after a real forward)
    assert "loss_moe" not in wrapped(None, None)                        # CPU scalars are the initial placeholder
    assert not list(wrapped.parameters()) or all(".moe." not in n_ for n_, _ in wrapped.named_parameters())
    with pytest.raises(ValueError):
        b200moe.registry.MoECriterion(Crit(), det, key="loss_aux_moe")


def test_context_dataset_subclass_reads_exported_fields(vendored):
    """The dataset hop (coco_dataset.py:43-62 drops image-level fields): a subclass built from a stand-in with the
    vendored class's constructor / load_item contract returns target['context']."""
    import b200moe
    from b200moe import context as cx
    ws, _, _ = vendored

    class CocoDetection:                                                # constructor + load_item of coco_dataset.py:28-61
        __inject__ = ["transforms"]
        __share__ = ["remap_mscoco_category"]

        def __init__(self, img_folder, ann_file, transforms, return_masks=False, remap_mscoco_category=False):
            self.ids = [7, 9]
            self.coco = types.SimpleNamespace(imgs={7: {"id": 7, "solar_context_bin": "high_sun(>45)", "road_type": "highway"},
                                                    9: {"id": 9}})

        def load_item(self, idx):
            return "img", {"boxes": torch.zeros(0, 4), "image_id": torch.tensor([self.ids[idx]])}

    cls = b200moe.registry.make_context_dataset(CocoDetection, name="ContextCocoDetectionT")
    ws.register()(cls)
    assert {"img_folder", "ann_file", "transforms"} <= set(ws.GLOBAL_CONFIG["ContextCocoDetectionT"]["_kwargs"])
    ds = cls("imgs", "ann.json", None)
    _, t0 = ds.load_item(0)
    _, t1 = ds.load_item(1)
    assert t0["context"].tolist() == cx.context_ids({"solar_context_bin": "high_sun(>45)", "road_type": "highway"})
    assert t0["context"].dtype == torch.long and t1["context"].tolist() == [0] * 5
