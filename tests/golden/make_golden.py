"""Generate tests/golden/*.npz from the UNMODIFIED vendored reference code.

Run once in the dev container (where /root/reference exists):
    python tests/golden/make_golden.py
The GPU box has no /root/reference; tests only read the committed .npz files.

Source of truth:  third_party/rtdetr/rtdetrv2_pytorch/src/zoo/rtdetr/hybrid_encoder.py
  * TransformerEncoderLayer (:116-161): the dense FFN at :157 is the op the MoE replaces
    (E identical experts + renormalised gates must reproduce it), and the post-norm
    residual wrapper at :154-160 is the step either side of the path.
The package __init__ chain needs faster_coco_eval (absent), so the three files are loaded
through empty stub packages.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/third_party/rtdetr/rtdetrv2_pytorch/src"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_vendored():
    for name in ("_rtv2", "_rtv2.core", "_rtv2.zoo", "_rtv2.zoo.rtdetr"):
        m = types.ModuleType(name)
        m.__path__ = []
        sys.modules[name] = m

    def load(modname, path):
        spec = importlib.util.spec_from_file_location(modname, path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[modname] = mod
        spec.loader.exec_module(mod)
        return mod

    ws = load("_rtv2.core.workspace", f"{REF}/core/workspace.py")
    sys.modules["_rtv2.core"].register = ws.register
    load("_rtv2.zoo.rtdetr.utils", f"{REF}/zoo/rtdetr/utils.py")
    return load("_rtv2.zoo.rtdetr.hybrid_encoder", f"{REF}/zoo/rtdetr/hybrid_encoder.py")


def main():
    he = load_vendored()
    torch.manual_seed(20260220)
    d, F, B, HW = 64, 128, 2, 25
    layer = he.TransformerEncoderLayer(d, 8, F, 0.0, "gelu")
    layer.eval()
    x = torch.randn(B, HW, d, requires_grad=True)
    go = torch.randn(B, HW, d)

    # the dense FFN exactly as hybrid_encoder.py:157 spells it
    y = layer.linear2(layer.dropout(layer.activation(layer.linear1(x))))
    (y * go).sum().backward()
    out = {
        "x": x.detach().numpy(), "go": go.numpy(), "y": y.detach().numpy(), "dx": x.grad.numpy(),
        "w1": layer.linear1.weight.detach().numpy(), "b1": layer.linear1.bias.detach().numpy(),
        "w2": layer.linear2.weight.detach().numpy(), "b2": layer.linear2.bias.detach().numpy(),
        "dw1": layer.linear1.weight.grad.numpy(), "db1": layer.linear1.bias.grad.numpy(),
        "dw2": layer.linear2.weight.grad.numpy(), "db2": layer.linear2.bias.grad.numpy(),
    }
    np.savez_compressed(os.path.join(HERE, "dense_ffn_d64_f128.npz"), **out)

    # same op at the smallest shape the tcgen05 tiles accept (d, F multiples of 128): the GPU known-answer test
    torch.manual_seed(20260221)
    d2, F2, B2, HW2 = 128, 256, 2, 50
    layer2 = he.TransformerEncoderLayer(d2, 8, F2, 0.0, "gelu")
    layer2.eval()
    x2 = torch.randn(B2, HW2, d2, requires_grad=True)
    go2 = torch.randn(B2, HW2, d2)
    y2 = layer2.linear2(layer2.dropout(layer2.activation(layer2.linear1(x2))))
    (y2 * go2).sum().backward()
    out2 = {
        "x": x2.detach().numpy(), "go": go2.numpy(), "y": y2.detach().numpy(), "dx": x2.grad.numpy(),
        "w1": layer2.linear1.weight.detach().numpy(), "b1": layer2.linear1.bias.detach().numpy(),
        "w2": layer2.linear2.weight.detach().numpy(), "b2": layer2.linear2.bias.detach().numpy(),
        "dw1": layer2.linear1.weight.grad.numpy(), "db1": layer2.linear1.bias.grad.numpy(),
        "dw2": layer2.linear2.weight.grad.numpy(), "db2": layer2.linear2.bias.grad.numpy(),
    }
    np.savez_compressed(os.path.join(HERE, "dense_ffn_d128_f256.npz"), **out2)

    # the post-norm wrapper (:154-160): norm2(residual + ffn(src)),
    # at the smallest shape the CUDA path accepts (d, F multiples of 128), forward and backward, and the
    # whole vendored encoder layer (:143-161: attention, both residuals, both norms) on the same weights: the in-situ
    # known answer for b200moe.MoEEncoderLayer.from_dense (fused post-norm path) on the GPU
    torch.manual_seed(20260222)
    B3, HW3 = 3, 70
    src3 = torch.randn(B3, HW3, d2, requires_grad=True)
    go3 = torch.randn(B3, HW3, d2)
    with torch.no_grad():
        layer2.norm2.weight.uniform_(0.5, 1.5)
        layer2.norm2.bias.uniform_(-0.2, 0.2)
    for p_ in layer2.parameters():
        p_.grad = None
    wrapped3 = layer2.norm2(src3 + layer2.dropout2(layer2.linear2(layer2.dropout(layer2.activation(layer2.linear1(src3))))))
    (wrapped3 * go3).sum().backward()
    np.savez_compressed(os.path.join(HERE, "postnorm_ffn_d128_f256.npz"), src=src3.detach().numpy(), go=go3.numpy(),
                        out=wrapped3.detach().numpy(), dsrc=src3.grad.numpy(),
                        dnorm2_weight=layer2.norm2.weight.grad.numpy(), dnorm2_bias=layer2.norm2.bias.grad.numpy(),
                        dw1=layer2.linear1.weight.grad.numpy(), dw2=layer2.linear2.weight.grad.numpy(),
                        **{"sd." + k: v.numpy() for k, v in layer2.state_dict().items()})
    for p_ in layer2.parameters():
        p_.grad = None
    pos3 = torch.randn(1, HW3, d2)
    xin = src3.detach().clone().requires_grad_(True)
    full3 = layer2(xin, pos_embed=pos3)
    (full3 * go3).sum().backward()
    np.savez_compressed(os.path.join(HERE, "encoder_layer_d128_f256.npz"), src=xin.detach().numpy(), pos=pos3.numpy(),
                        go=go3.numpy(), out=full3.detach().numpy(), dsrc=xin.grad.numpy(),
                        **{"sd." + k: v.numpy() for k, v in layer2.state_dict().items()})
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
