"""Binding to the vendored plugin registry (SURVEY.md §8(b) "How the host picks it up" (i)).

The RT-DETRv2 stack builds every module from YAML through `@register()` / `create()`
(third_party/rtdetr/rtdetrv2_pytorch/src/core/workspace.py:15-53, 92-179): `register` records the constructor's
*named* arguments and their defaults (`inspect.getfullargspec`, :63-87 — a `**kwargs` does not survive that), `create`
calls the class with the merged YAML values and resolves it as an attribute of its defining module (:110).
`HybridEncoder` itself hard-codes its layer class (hybrid_encoder.py:232-241), so the MoE enters the YAML as an encoder
class of its own: `MoEHybridEncoder`, a subclass of the vendored `HybridEncoder` with the same constructor arguments
(:186-201) plus the MoE ones, which upcycles the AIFI layers it has just built.

The vendored classes are not importable from this package (and must not be copied), so the subclasses are created by
factories that receive them:

    # e.g. at the top of third_party/rtdetr/rtdetrv2_pytorch/tools/train.py, after `from src.core import ...`
    import b200moe
    from src.core import register
    from src.zoo.rtdetr.hybrid_encoder import HybridEncoder
    from src.data.dataset.coco_dataset import CocoDetection
    b200moe.registry.register_all(register, HybridEncoder, CocoDetection)

    # configs/rtdetrv2/include/rtdetrv2_r50vd.yml
    RTDETR:
      encoder: MoEHybridEncoder
    MoEHybridEncoder:
      in_channels: [128, 256, 512]
      hidden_dim: 256
      expansion: 0.5
      num_experts: 8
      top_k: 2
      capacity_factor: 1.25
"""
from __future__ import annotations

import sys
from typing import Optional, Sequence

import torch

from .context import CONTEXT_FIELDS, CONTEXT_VOCAB_SIZES, attach_context
from .encoder import MoEEncoderLayer, upcycle_encoder
from .layer import ContextMoE

_THIS = sys.modules[__name__]


def _publish(cls, name: str):
    """`create()` looks the class up as `getattr(importlib.import_module(cls.__module__), name)` (workspace.py:83,110)."""
    cls.__name__ = cls.__qualname__ = name
    cls.__module__ = __name__
    setattr(_THIS, name, cls)
    return cls


def make_moe_hybrid_encoder(HybridEncoder, name: str = "MoEHybridEncoder"):
    """Subclass of the vendored HybridEncoder whose AIFI layers carry a ContextMoE instead of the dense FFN."""

    class _MoEHybridEncoder(HybridEncoder):
        __share__ = list(getattr(HybridEncoder, "__share__", []))

        # the vendored constructor's arguments and defaults (hybrid_encoder.py:186-201), then the MoE's
        def __init__(self, in_channels=[512, 1024, 2048], feat_strides=[8, 16, 32], hidden_dim=256, nhead=8,  # noqa: B006
                     dim_feedforward=1024, dropout=0.0, enc_act="gelu", use_encoder_idx=[2], num_encoder_layers=1,  # noqa: B006
                     pe_temperature=10000, expansion=1.0, depth_mult=1.0, act="silu", eval_spatial_size=None,
                     version="v2", num_experts=8, top_k=2, capacity_factor=1.25,
                     context_vocab_sizes=CONTEXT_VOCAB_SIZES, context_dim=32, aux_loss_weight=1e-2,
                     z_loss_weight=1e-3, normalize_topk=True):
            super().__init__(in_channels=in_channels, feat_strides=feat_strides, hidden_dim=hidden_dim, nhead=nhead,
                             dim_feedforward=dim_feedforward, dropout=dropout, enc_act=enc_act,
                             use_encoder_idx=use_encoder_idx, num_encoder_layers=num_encoder_layers,
                             pe_temperature=pe_temperature, expansion=expansion, depth_mult=depth_mult, act=act,
                             eval_spatial_size=eval_spatial_size, version=version)
            self.num_moe_layers = upcycle_encoder(
                self, num_experts=num_experts, top_k=top_k, capacity_factor=capacity_factor,
                context_vocab_sizes=tuple(context_vocab_sizes), context_dim=context_dim,
                aux_loss_weight=aux_loss_weight, z_loss_weight=z_loss_weight, normalize_topk=normalize_topk,
                activation=enc_act)

    return _publish(_MoEHybridEncoder, name)


def make_context_dataset(CocoDetection, name: str = "ContextCocoDetection",
                         fields: Sequence[str] = CONTEXT_FIELDS):
    """Subclass of the vendored CocoDetection whose targets carry `context` (LongTensor [n_fields]): the image-level
    fields scripts/export_coco_dataset.py writes into `images[]` (:145-148) are read back from `self.coco.imgs` instead
    of being dropped at coco_dataset.py:43-46, 156-171.  `det_engine.py:40` moves the tensor to the device with the
    boxes; `b200moe.install_context_hook(model)` hands it to the encoder."""

    class _ContextCocoDetection(CocoDetection):
        __inject__ = list(getattr(CocoDetection, "__inject__", []))
        __share__ = list(getattr(CocoDetection, "__share__", []))

        def __init__(self, img_folder, ann_file, transforms, return_masks=False, remap_mscoco_category=False):
            super().__init__(img_folder, ann_file, transforms, return_masks, remap_mscoco_category)

        def load_item(self, idx):
            image, target = super().load_item(idx)
            record = self.coco.imgs.get(self.ids[idx], {}) if hasattr(self, "coco") else {}
            attach_context(target, record, fields)
            return image, target

    return _publish(_ContextCocoDetection, name)


def register_all(register, HybridEncoder=None, CocoDetection=None) -> dict:
    """Register the layer classes (explicit constructor arguments, as the schema extraction needs them) and, when the
    vendored classes are passed, the YAML-facing subclasses.  Returns {name: class}."""
    out = {"ContextMoE": ContextMoE, "MoEEncoderLayer": MoEEncoderLayer}
    if HybridEncoder is not None:
        out["MoEHybridEncoder"] = make_moe_hybrid_encoder(HybridEncoder)
    if CocoDetection is not None:
        out["ContextCocoDetection"] = make_context_dataset(CocoDetection)
    for cls in out.values():
        register()(cls)
    return out


class MoECriterion(torch.nn.Module):
    """Wraps a detection criterion so that the MoE auxiliary loss reaches `sum(loss_dict.values())`
    (det_engine.py:48-51) without editing the engine: the key is added AFTER the criterion's weight_dict filter
    (rtdetrv2_criterion.py:164) and does not contain 'aux' (:146 treats such keys as auxiliary decoder outputs)."""

    def __init__(self, criterion: torch.nn.Module, model: torch.nn.Module, key: str = "loss_moe"):
        super().__init__()
        if "aux" in key:
            raise ValueError("the loss key must not contain 'aux' (rtdetrv2_criterion.py:146)")
        self.criterion, self.key = criterion, key
        self.__dict__["_model"] = model            # not a submodule: the criterion must not own the model's parameters

    def forward(self, outputs, targets, **kwargs):
        from .context import collect_aux_loss
        loss_dict = self.criterion(outputs, targets, **kwargs)
        aux: Optional[torch.Tensor] = collect_aux_loss(self.__dict__["_model"])
        if aux is not None:
            loss_dict[self.key] = aux
        return loss_dict
