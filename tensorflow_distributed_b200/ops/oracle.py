"""Plain-PyTorch fp32 restatement of every numeric op the reference calls.

This is the correctness oracle for the sm_100a kernels and the compute path of
the CPU configs.  TensorFlow 1.x semantics are kept: NHWC activations, HWIO
filters, SAME padding, inverted dropout, dense-label softmax cross entropy,
TF-form Adam (epsilon outside the bias-corrected sqrt).  Call sites in the
reference: ``/root/reference/mnist_python_m.py:93-128`` (model), ``:205-208``
(loss / accuracy / Adam).
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F


def conv2d_same_nhwc(x: torch.Tensor, w_hwio: torch.Tensor) -> torch.Tensor:
    """tf.nn.conv2d(strides=1, padding='SAME') on NHWC input with an HWIO filter (m:95)."""
    kh, kw = w_hwio.shape[0], w_hwio.shape[1]
    y = F.conv2d(x.permute(0, 3, 1, 2), w_hwio.permute(3, 2, 0, 1), padding=(kh // 2, kw // 2))
    return y.permute(0, 2, 3, 1)


def maxpool2x2_nhwc(x: torch.Tensor) -> torch.Tensor:
    """tf.nn.max_pool(ksize=2, strides=2, 'SAME') (m:101); dims are even here so SAME == VALID."""
    return F.max_pool2d(x.permute(0, 3, 1, 2), 2, 2, ceil_mode=True).permute(0, 2, 3, 1)


def dropout_with_mask(x: torch.Tensor, keep_mask: Optional[torch.Tensor], keep_prob: float) -> torch.Tensor:
    """tf.nn.dropout: kept units scaled by 1/keep_prob (m:124)."""
    if keep_mask is None or keep_prob >= 1.0:
        return x
    return x * keep_mask.to(x.dtype) / keep_prob


def softmax_xent_mean(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """reduce_mean(softmax_cross_entropy_with_logits(logits, dense labels)) (m:205)."""
    return -(labels * F.log_softmax(logits, dim=1)).sum(dim=1).mean()


def accuracy(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """mean(argmax(pred)==argmax(y)) (m:206-207)."""
    return (logits.argmax(dim=1) == labels.argmax(dim=1)).float().mean()


def quantizer(precision: Optional[str]):
    """Operand quantisation of a tensor-core precision arm as a straight-through function (identity gradient):
    'bf16' rounds to bfloat16, 'tf32' drops the low 13 mantissa bits (what tcgen05 kind::tf32 reads of an fp32 word),
    None / 'fp32' is the identity.  Used to derive test tolerances: oracle(quantised operands) - oracle(fp32) is the
    error a precision arm is ENTITLED to; what a kernel adds on top of it is accumulation order only."""
    if precision in (None, "fp32"):
        return lambda t: t
    if precision == "bf16":
        rnd = lambda t: t.to(torch.bfloat16).to(t.dtype)  # noqa: E731
    elif precision == "tf32":
        rnd = lambda t: (t.contiguous().view(torch.int32) & ~0x1fff).view(torch.float32)  # noqa: E731
    else:
        raise ValueError("unknown precision %r" % precision)
    return lambda t: t + (rnd(t.detach()) - t.detach())


def cnn_forward(x: torch.Tensor, p: Dict[str, torch.Tensor], keep_prob: float = 1.0,
                keep_mask: Optional[torch.Tensor] = None, q=None) -> torch.Tensor:
    """conv_net (m:104-128).  ``x``: [B,784]; returns logits [B,10].  ``q``: operand quantiser of the tensor-core
    layers (conv2, fc1); conv1 and the 10-wide output layer run on CUDA cores in fp32 in every arm."""
    q = q or (lambda t: t)
    h = x.reshape(-1, 28, 28, 1)
    h = torch.relu(conv2d_same_nhwc(h, p["weights/wc1"]) + p["biases/bc1"])
    h = maxpool2x2_nhwc(h)
    h = torch.relu(conv2d_same_nhwc(q(h), q(p["weights/wc2"])) + p["biases/bc2"])
    h = maxpool2x2_nhwc(h)
    h = h.reshape(-1, p["weights/wd1"].shape[0])          # NHWC flatten order (m:120)
    h = torch.relu(q(h) @ q(p["weights/wd1"]) + p["biases/bd1"])
    h = dropout_with_mask(h, keep_mask, keep_prob)
    return h @ p["weights/out"] + p["biases/out"]


def mlp_forward(x: torch.Tensor, p: Dict[str, torch.Tensor], keep_prob: float = 1.0,
                keep_mask: Optional[torch.Tensor] = None, q=None) -> torch.Tensor:
    """784 -> hidden x L (ReLU) -> 10.  Dropout is not part of the MLP."""
    q = q or (lambda t: t)
    h = x
    i = 1
    while "weights/h%d" % i in p:
        h = torch.relu(q(h) @ q(p["weights/h%d" % i]) + p["biases/b%d" % i])
        i += 1
    return h @ p["weights/out"] + p["biases/out"]


def forward(kind: str, x, p, keep_prob: float = 1.0, keep_mask=None, q=None):
    return (cnn_forward if kind == "cnn" else mlp_forward)(x, p, keep_prob, keep_mask, q)


def dropout_width(kind: str, p: Dict[str, torch.Tensor]) -> int:
    return p["weights/wd1"].shape[1] if kind == "cnn" else 0


def loss_and_grads(kind: str, x: torch.Tensor, y: torch.Tensor, p: Dict[str, torch.Tensor],
                   keep_prob: float = 1.0, keep_mask: Optional[torch.Tensor] = None, precision: Optional[str] = None
                   ) -> Tuple[torch.Tensor, Dict[str, torch.Tensor], torch.Tensor]:
    """Loss, d(loss)/d(param) for every parameter, and logits (autograd, fp32; ``precision`` quantises the operands
    of the tensor-core layers the way that arm's kernels see them)."""
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in p.items()}
    logits = forward(kind, x, leaves, keep_prob, keep_mask, quantizer(precision))
    loss = softmax_xent_mean(logits, y)
    grads = torch.autograd.grad(loss, [leaves[k] for k in leaves])
    return loss.detach(), {k: g for k, g in zip(leaves, grads)}, logits.detach()


class TFAdam:
    """tf.train.AdamOptimizer (m:208): beta1=0.9, beta2=0.999, eps=1e-8.

    ``lr_t = lr*sqrt(1-b2^t)/(1-b1^t); m = b1*m+(1-b1)*g; v = b2*v+(1-b2)*g*g;
    var -= lr_t*m/(sqrt(v)+eps)`` -- epsilon is added outside the bias
    correction, unlike torch.optim.Adam.
    """

    def __init__(self, lr: float = 0.01, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8):
        self.lr, self.beta1, self.beta2, self.eps = lr, beta1, beta2, eps
        self.t = 0
        self.m: Dict[str, torch.Tensor] = {}
        self.v: Dict[str, torch.Tensor] = {}

    def lr_t(self, t: int) -> float:
        return self.lr * math.sqrt(1.0 - self.beta2 ** t) / (1.0 - self.beta1 ** t)

    @torch.no_grad()
    def apply(self, params: Dict[str, torch.Tensor], grads: Dict[str, torch.Tensor]) -> None:
        self.t += 1
        lr_t = self.lr_t(self.t)
        for k, g in grads.items():
            if k not in self.m:
                self.m[k] = torch.zeros_like(params[k])
                self.v[k] = torch.zeros_like(params[k])
            m, v = self.m[k], self.v[k]
            m.mul_(self.beta1).add_(g, alpha=1.0 - self.beta1)
            v.mul_(self.beta2).addcmul_(g, g, value=1.0 - self.beta2)
            params[k].sub_(lr_t * m / (v.sqrt() + self.eps))

    def state_dict(self) -> Dict[str, object]:
        return {"t": self.t, "m": {k: v.clone() for k, v in self.m.items()},
                "v": {k: v.clone() for k, v in self.v.items()},
                "beta1_power": self.beta1 ** self.t, "beta2_power": self.beta2 ** self.t}

    def load_state_dict(self, sd: Dict[str, object]) -> None:
        self.t = int(sd["t"])
        self.m = {k: v.clone() for k, v in sd["m"].items()}
        self.v = {k: v.clone() for k, v in sd["v"].items()}


def init_params(spec, seed: int = 0, stddev: float = 1.0, device="cpu") -> Dict[str, torch.Tensor]:
    """tf.random_normal init, mean 0 stddev 1 for every parameter (m:186-195)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    return {v.name: (torch.randn(v.shape, generator=g, dtype=torch.float32) * stddev).to(device)
            for v in spec.variables}
