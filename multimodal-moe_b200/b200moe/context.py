"""Scene-context vocabulary and wire-through.

The reference computes the context per image offline and exports part of it into the COCO json, then drops it when
the dataset is loaded:
  * solar elevation -> 5 bins + "missing": scripts/add_solar_context_bins.py:90-107
  * categorical ZOD fields and their category lists: scripts/analyze_context_frequencies.py:60-66,89 and the fixture
    outputs/analysis/camera/detection/context_field_frequencies_final.csv:2-26
  * export into `images[]`: scripts/export_coco_dataset.py:145-148,246-254
  * where it is lost: third_party/rtdetr/rtdetrv2_pytorch/src/data/dataset/coco_dataset.py:43-62 (target dict built
    without the image-level fields) and src/zoo/rtdetr/rtdetr.py:32-37 (targets reach the decoder only)

This module turns the exported strings into the integer ids `ContextMoE.forward(x, context)` takes (one id per field
per image, id 0 = "missing", remaining ids in the order of `solar_labels` for the solar field and sorted category name
for the categorical fields — Appendix A.2), and carries them from the detection targets to every ContextMoE of a
model without editing the training engine (`det_engine.py:40` moves LongTensors in `targets` to the device already).
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Mapping, Optional, Sequence

import torch
import torch.nn as nn

CONTEXT_FIELDS = ("solar_context_bin", "scraped_weather", "time_of_day", "road_type", "road_condition")
CONTEXT_VOCAB: Dict[str, List[str]] = {
    "solar_context_bin": ["missing", "night(<-6)", "twilight(-6..0)", "low_sun(0..15)", "mid_sun(15..45)",
                          "high_sun(>45)"],
    "scraped_weather": ["missing"] + sorted(["partly-cloudy-day", "cloudy", "clear-day", "rain", "partly-cloudy-night",
                                             "clear-night", "fog", "snow", "wind"]),
    "time_of_day": ["missing"] + sorted(["day", "night", "twilight"]),
    "road_type": ["missing"] + sorted(["city", "arterial-urban", "highway", "arterial-rural", "smaller-rural"]),
    "road_condition": ["missing"] + sorted(["normal", "wet", "snow"]),
}
CONTEXT_VOCAB_SIZES = tuple(len(CONTEXT_VOCAB[f]) for f in CONTEXT_FIELDS)   # (6, 10, 4, 6, 4)
_LOOKUP = {f: {name: i for i, name in enumerate(CONTEXT_VOCAB[f])} for f in CONTEXT_FIELDS}


def context_ids(image_record: Mapping, fields: Sequence[str] = CONTEXT_FIELDS) -> List[int]:
    """COCO `images[]` record (or any mapping with the exported string fields) -> one id per field.
    Absent fields, None, NaN and unknown categories map to 0 ("missing"), like add_solar_context_bins.py:107."""
    out = []
    for f in fields:
        v = image_record.get(f)
        if v is None or (isinstance(v, float) and v != v):
            out.append(0)
        else:
            out.append(_LOOKUP[f].get(str(v), 0))
    return out


def attach_context(target: dict, image_record: Mapping, fields: Sequence[str] = CONTEXT_FIELDS) -> dict:
    """What a maintainer adds at coco_dataset.py:62 (after the target dict is built): target['context'] =
    LongTensor[n_fields].  A LongTensor, so det_engine.py:40's `.to(device)` moves it with the boxes."""
    target["context"] = torch.tensor(context_ids(image_record, fields), dtype=torch.long)
    return target


def batch_context(targets: Optional[Iterable[Mapping]], n_fields: int = len(CONTEXT_FIELDS)) -> Optional[torch.Tensor]:
    """targets (list of per-image dicts, det_engine.py:38-41) -> LongTensor [B, n_fields]; None when no target carries a
    context (evaluation without metadata: the layer then routes on the tokens alone)."""
    if targets is None:
        return None
    rows = []
    for t in targets:
        c = t.get("context") if isinstance(t, Mapping) else None
        if c is None:
            return None
        rows.append(torch.as_tensor(c, dtype=torch.long).reshape(n_fields))
    return torch.stack(rows) if rows else None


def moe_layers(model: nn.Module):
    from .layer import ContextMoE
    return [m for m in model.modules() if isinstance(m, ContextMoE)]


def install_context_hook(model: nn.Module) -> torch.utils.hooks.RemovableHandle:
    """Forward pre-hook on the detector (`RTDETR.forward(x, targets=None)`, rtdetr.py:32): stores the batch's context on
    every module that has a `set_context` method (MoEEncoderLayer) before the encoder runs.  Returns the hook handle."""
    def hook(mod, args, kwargs):
        targets = kwargs.get("targets", args[1] if len(args) > 1 else None)
        ctx = batch_context(targets)
        for m in mod.modules():
            if hasattr(m, "set_context"):
                m.set_context(ctx)
        return None
    return model.register_forward_pre_hook(hook, with_kwargs=True)


def collect_aux_loss(model: nn.Module) -> Optional[torch.Tensor]:
    """Sum of the weighted auxiliary losses of every ContextMoE after a forward (det_engine.py:48-51: add it to
    loss_dict under a key without 'aux', after the criterion's weight_dict filter, rtdetrv2_criterion.py:146,164)."""
    losses = [m.aux_loss for m in moe_layers(model) if isinstance(m.aux_loss, torch.Tensor) and m.aux_loss.is_cuda]
    return torch.stack([l_.float() for l_ in losses]).sum() if losses else None


def router_telemetry(model: nn.Module) -> Dict[str, float]:
    """Per-layer gating telemetry for log.txt / TensorBoard (the proposal asks for it:
    notes/MoE_in_ZOD_Thesis_Proposal_revisedTimeline.txt:152,240).  One host sync; call it at logging frequency."""
    out: Dict[str, float] = {}
    for i, m in enumerate(moe_layers(model)):
        a = m.last_aux
        if not a:
            continue
        tpe = a["tokens_per_expert"].float()
        out[f"moe{i}/load_balance"] = float(a["load_balance"])
        out[f"moe{i}/z_loss"] = float(a["z_loss"])
        out[f"moe{i}/dropped"] = float(a["dropped"])
        out[f"moe{i}/max_expert_share"] = float(tpe.max() / tpe.sum().clamp_min(1))
    return out
