"""Interop: trained parameters as a stock ``torch.nn.Module`` (NCHW convolutions, ``nn.Linear``).

Checkpoints and engines keep the reference's layouts -- NHWC activations, HWIO filters, ``[in, out]`` dense weights,
flatten in H,W,C order (``/root/reference/mnist_python_m.py:106-127``).  PyTorch's defaults differ in every one of
these, so a hand conversion is easy to get subtly wrong (the rows of ``wd1`` have to be permuted from h,w,c to
c,h,w order).  ``to_torch_module`` does it once, and ``from_torch_module`` goes the other way, so a model trained
here can be served by ordinary PyTorch code and a PyTorch-trained one can be loaded into the engines.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict

import torch
from torch import nn

from .models.spec import ModelSpec


class _Flatten784(nn.Module):
    """Accept [B,784], [B,28,28], [B,28,28,1] or [B,1,28,28] like the placeholder reshape of the reference (m:106)."""

    def __init__(self, image: bool):
        super().__init__()
        self.image = image

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x.reshape(x.shape[0], -1)
        return x.reshape(-1, 1, 28, 28) if self.image else x


def to_torch_module(spec: ModelSpec, params: Dict[str, torch.Tensor]) -> nn.Module:
    """Inference module (dropout off, as with keep_prob = 1) whose logits equal the engines' on the same input."""
    p = {k: v.detach().to(torch.float32).cpu() for k, v in params.items()}
    if spec.kind == "cnn":
        conv1, conv2 = nn.Conv2d(1, 32, 5, padding=2), nn.Conv2d(32, 64, 5, padding=2)
        fc1, out = nn.Linear(7 * 7 * 64, 1024), nn.Linear(1024, 10)
        with torch.no_grad():
            conv1.weight.copy_(p["weights/wc1"].permute(3, 2, 0, 1))   # HWIO -> OIHW
            conv2.weight.copy_(p["weights/wc2"].permute(3, 2, 0, 1))
            conv1.bias.copy_(p["biases/bc1"]); conv2.bias.copy_(p["biases/bc2"])
            # wd1 rows are indexed h*448 + w*64 + c (NHWC flatten); nn.Flatten on NCHW gives c*49 + h*7 + w
            wd1 = p["weights/wd1"].reshape(7, 7, 64, 1024).permute(2, 0, 1, 3).reshape(3136, 1024)
            fc1.weight.copy_(wd1.t()); fc1.bias.copy_(p["biases/bd1"])
            out.weight.copy_(p["weights/out"].t()); out.bias.copy_(p["biases/out"])
        return nn.Sequential(OrderedDict([
            ("input", _Flatten784(True)), ("conv1", conv1), ("relu1", nn.ReLU()), ("pool1", nn.MaxPool2d(2)),
            ("conv2", conv2), ("relu2", nn.ReLU()), ("pool2", nn.MaxPool2d(2)), ("flatten", nn.Flatten()),
            ("fc1", fc1), ("relu3", nn.ReLU()), ("out", out)])).eval()
    layers = [("input", _Flatten784(False))]
    dims = [784] + list(spec.hidden)
    for i in range(1, len(dims)):
        lin = nn.Linear(dims[i - 1], dims[i])
        with torch.no_grad():
            lin.weight.copy_(p["weights/h%d" % i].t()); lin.bias.copy_(p["biases/b%d" % i])
        layers += [("h%d" % i, lin), ("relu%d" % i, nn.ReLU())]
    out = nn.Linear(dims[-1], 10)
    with torch.no_grad():
        out.weight.copy_(p["weights/out"].t()); out.bias.copy_(p["biases/out"])
    layers.append(("out", out))
    return nn.Sequential(OrderedDict(layers)).eval()


def from_torch_module(spec: ModelSpec, module: nn.Module) -> Dict[str, torch.Tensor]:
    """Inverse of ``to_torch_module``: parameters in the reference's names and layouts (for ``engine.set_params``)."""
    sd = {k: v.detach().to(torch.float32).cpu() for k, v in module.state_dict().items()}
    if spec.kind == "cnn":
        wd1 = sd["fc1.weight"].t().reshape(64, 7, 7, 1024).permute(1, 2, 0, 3).reshape(3136, 1024)
        return {"weights/wc1": sd["conv1.weight"].permute(2, 3, 1, 0).contiguous(),
                "weights/wc2": sd["conv2.weight"].permute(2, 3, 1, 0).contiguous(),
                "weights/wd1": wd1.contiguous(), "weights/out": sd["out.weight"].t().contiguous(),
                "biases/bc1": sd["conv1.bias"], "biases/bc2": sd["conv2.bias"], "biases/bd1": sd["fc1.bias"],
                "biases/out": sd["out.bias"]}
    out = {"weights/out": sd["out.weight"].t().contiguous(), "biases/out": sd["out.bias"]}
    for i in range(1, len(spec.hidden) + 1):
        out["weights/h%d" % i] = sd["h%d.weight" % i].t().contiguous()
        out["biases/b%d" % i] = sd["h%d.bias" % i]
    return out
