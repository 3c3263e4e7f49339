"""Encoder-layer host glue: the step either side of the path.

`MoEEncoderLayer` has the constructor and forward signature of the vendored
`TransformerEncoderLayer(d_model, nhead, dim_feedforward, dropout, activation, normalize_before)`
(third_party/rtdetr/rtdetrv2_pytorch/src/zoo/rtdetr/hybrid_encoder.py:116-161) with the dense FFN of :129-131/:157
replaced by `ContextMoE`; `HybridEncoder` builds its AIFI stack from such a layer (:238-241, deep-copied at :167), so
`upcycle_encoder` can swap the layers of a constructed (or checkpoint-loaded) model in place: attention and norms
are kept, every expert starts as a copy of the dense FFN, so with renormalised gates and no drops the swapped model
computes what the dense one did (the K1 anchor of SURVEY.md §8(c)) and training departs from there.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.nn as nn

from .layer import ContextMoE


class MoEEncoderLayer(nn.Module):
    # every argument is named: the vendored registry records the constructor's argspec (workspace.py:63-87), a
    # **kwargs would be dropped by `create()`
    def __init__(self, d_model: int, nhead: int, dim_feedforward: int = 2048, dropout: float = 0.0,
                 activation: str = "gelu", normalize_before: bool = False, num_experts: int = 8, top_k: int = 2,
                 capacity_factor: Optional[float] = 1.25, context_vocab_sizes: Sequence[int] = (),
                 context_dim: int = 32, aux_loss_weight: float = 1e-2, z_loss_weight: float = 1e-3,
                 normalize_topk: bool = True, ep_group=None):
        super().__init__()
        if dropout > 0.0:
            # hybrid_encoder.py:157 applies self.dropout between the activation and linear2; the expert FFNs have no
            # inner dropout (A.6: dropout 0, as rtdetrv2_r50vd.yml:38 configures the encoder)
            raise ValueError("MoEEncoderLayer supports dropout=0.0 only (the vendored encoder default)")
        self.normalize_before = normalize_before
        self.self_attn = nn.MultiheadAttention(d_model, nhead, dropout, batch_first=True)
        self.moe = ContextMoE(d_model, dim_feedforward, num_experts, top_k=top_k, capacity_factor=capacity_factor,
                              context_vocab_sizes=context_vocab_sizes, context_dim=context_dim, activation=activation,
                              aux_loss_weight=aux_loss_weight, z_loss_weight=z_loss_weight,
                              normalize_topk=normalize_topk, ep_group=ep_group)
        self.norm1 = nn.LayerNorm(d_model)
        self.norm2 = nn.LayerNorm(d_model)
        self.dropout1 = nn.Dropout(dropout)
        self.dropout2 = nn.Dropout(dropout)
        self._context: Optional[torch.Tensor] = None

    def set_context(self, context: Optional[torch.Tensor]) -> None:
        """LongTensor [B, n_fields] of the batch about to be encoded (b200moe.context.install_context_hook), or None."""
        self._context = context

    @property
    def aux_loss(self) -> torch.Tensor:
        return self.moe.aux_loss

    def forward(self, src: torch.Tensor, src_mask=None, pos_embed=None) -> torch.Tensor:
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
        ctx = self._context
        if ctx is not None and ctx.device != src.device:
            ctx = ctx.to(src.device)
        if not self.normalize_before and (self.dropout2.p == 0.0 or not self.training):
            # post-norm (the vendored default): residual add + LayerNorm ride the combine kernel (:157-160 in one pass)
            return self.moe.forward_post_norm(src, residual, self.norm2, ctx)
        src = self.moe(src, ctx)                     # hybrid_encoder.py:157; dropped tokens give 0, the residual carries them
        src = residual + self.dropout2(src)
        if not self.normalize_before:
            src = self.norm2(src)
        return src

    @classmethod
    def from_dense(cls, dense: nn.Module, num_experts: int, **moe_kwargs) -> "MoEEncoderLayer":
        """Build from a constructed dense layer (anything with self_attn, linear1, linear2, norm1, norm2, dropout1,
        dropout2, normalize_before): shares its attention / norm modules, copies the dense FFN into every expert."""
        d_model, dff = dense.linear1.in_features, dense.linear1.out_features
        inner = getattr(dense, "dropout", None)
        if isinstance(inner, nn.Dropout) and inner.p > 0.0:
            raise ValueError("the dense layer applies dropout inside its FFN (hybrid_encoder.py:157); the expert FFNs "
                             "have none (A.6) -- build the encoder with dropout=0.0 (rtdetrv2_r50vd.yml:38)")
        if moe_kwargs.get("activation") is None:
            act = getattr(dense, "activation", None)          # zoo/rtdetr/utils.py:137-172 builds these modules
            if act is None or isinstance(act, nn.GELU) or act is torch.nn.functional.gelu:
                moe_kwargs["activation"] = "gelu"
            elif isinstance(act, nn.ReLU) or act is torch.nn.functional.relu:
                moe_kwargs["activation"] = "relu"
            else:
                raise ValueError(f"unsupported FFN activation {act!r}: the expert epilogues implement gelu and relu")
        new = cls.__new__(cls)
        nn.Module.__init__(new)
        new.normalize_before = bool(getattr(dense, "normalize_before", False))
        new.self_attn, new.norm1, new.norm2 = dense.self_attn, dense.norm1, dense.norm2
        new.dropout1, new.dropout2 = dense.dropout1, dense.dropout2
        new.moe = ContextMoE(d_model, dff, num_experts, **moe_kwargs)
        new._context = None
        with torch.no_grad():
            new.moe.w1.copy_(dense.linear1.weight.detach().float().expand_as(new.moe.w1))
            new.moe.b1.copy_(dense.linear1.bias.detach().float().expand_as(new.moe.b1))
            new.moe.w2.copy_(dense.linear2.weight.detach().float().expand_as(new.moe.w2))
            new.moe.b2.copy_(dense.linear2.bias.detach().float().expand_as(new.moe.b2))
        return new.to(dense.linear1.weight.device)


def upcycle_encoder(model: nn.Module, num_experts: int, layer_type_name: str = "TransformerEncoderLayer",
                    **moe_kwargs) -> int:
    """Replace every dense encoder layer of `model` (matched by class name, so that the vendored class does not have
    to be importable here) by a MoEEncoderLayer built with `from_dense`.  Returns the number of layers swapped."""
    n = 0
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if type(child).__name__ == layer_type_name and hasattr(child, "linear1") and hasattr(child, "self_attn"):
                setattr(parent, name, MoEEncoderLayer.from_dense(child, num_experts, **moe_kwargs))
                n += 1
    return n
