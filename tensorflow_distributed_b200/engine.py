"""Execution engines: what the reference delegates to ``tf.Session.run``.

``OracleEngine`` runs the plain-PyTorch fp32 restatement (CPU configs, and the
oracle the GPU tests compare against).  ``NativeEngine`` (``engine_native.py``)
runs the hand-written sm_100a kernels with an explicit backward pass.  Both
expose the three fetches the reference ever asks a session for: the train op
(``/root/reference/mnist_single.py:112``), the loss with keep_prob=1
(``single:115``) and accuracy (``single:129``, ``mnist_python_m.py:315``).
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

from .models.spec import ModelSpec
from .ops import oracle


def _as_tensor(a, device="cpu") -> torch.Tensor:
    if isinstance(a, np.ndarray):
        a = torch.from_numpy(np.ascontiguousarray(a))
    return a.to(device=device, dtype=torch.float32)


class OracleEngine:
    """Single-process engine: parameters, TF-Adam and autograd gradients on one torch device."""

    name = "oracle"

    def __init__(self, spec: ModelSpec, learning_rate: float = 0.01, seed: int = 0, init_stddev: float = 1.0,
                 device: str = "cpu"):
        self.spec, self.device = spec, device
        self.params: Dict[str, torch.Tensor] = oracle.init_params(spec, seed, init_stddev, device)
        self.adam = oracle.TFAdam(learning_rate)
        self.global_step = 0
        self._gen = torch.Generator(device="cpu").manual_seed(seed + 7919)

    # -- gradients only (used by cluster workers: the apply happens on the ps) --
    def draw_keep_mask(self, batch: int, keep_prob: float) -> Optional[torch.Tensor]:
        width = oracle.dropout_width(self.spec.kind, self.params)
        if keep_prob >= 1.0 or width == 0:
            return None
        return (torch.rand(batch, width, generator=self._gen) < keep_prob).to(self.device)

    def loss_and_grads(self, x, y, keep_prob: float = 1.0, params=None, keep_mask=None):
        x, y = _as_tensor(x, self.device), _as_tensor(y, self.device)
        if keep_mask is None:
            keep_mask = self.draw_keep_mask(x.shape[0], keep_prob)
        return oracle.loss_and_grads(self.spec.kind, x, y, params or self.params, keep_prob, keep_mask)

    # -- the three session fetches --
    def train_step(self, x, y, keep_prob: float = 0.75) -> float:
        loss, grads, _ = self.loss_and_grads(x, y, keep_prob)
        self.adam.apply(self.params, grads)
        self.global_step += 1
        return float(loss)

    @torch.no_grad()
    def eval_loss(self, x, y) -> float:
        x, y = _as_tensor(x, self.device), _as_tensor(y, self.device)
        return float(oracle.softmax_xent_mean(oracle.forward(self.spec.kind, x, self.params), y))

    @torch.no_grad()
    def eval_accuracy(self, x, y) -> float:
        x, y = _as_tensor(x, self.device), _as_tensor(y, self.device)
        return float(oracle.accuracy(oracle.forward(self.spec.kind, x, self.params), y))

    def predict_proba(self, x):
        """Serving call: class probabilities softmax(model(x)) for a host batch, [B, 10] float32 (dropout off)."""
        x = _as_tensor(x, self.device).reshape(-1, 784)
        with torch.no_grad():
            return torch.softmax(oracle.forward(self.spec.kind, x, self.params), dim=1).cpu().numpy()

    def predict(self, x):
        """argmax class per image (``tf.argmax(pred, 1)``, m:206)."""
        return self.predict_proba(x).argmax(axis=1)

    # -- checkpoint plumbing --
    def get_params(self) -> Dict[str, torch.Tensor]:
        return {k: v.detach().cpu().clone() for k, v in self.params.items()}

    def set_params(self, params: Dict[str, torch.Tensor]) -> None:
        self.params = {k: v.to(self.device, torch.float32).clone() for k, v in params.items()}

    def get_adam_state(self) -> dict:
        sd = self.adam.state_dict()
        for k in self.params:
            sd["m"].setdefault(k, torch.zeros_like(self.params[k]))
            sd["v"].setdefault(k, torch.zeros_like(self.params[k]))
        return sd

    def set_adam_state(self, sd: dict) -> None:
        self.adam.load_state_dict({"t": sd["t"], "m": {k: v.to(self.device) for k, v in sd["m"].items()},
                                   "v": {k: v.to(self.device) for k, v in sd["v"].items()}})

    def synchronize(self) -> None:
        if str(self.device).startswith("cuda"):
            torch.cuda.synchronize()


def make_engine(spec: ModelSpec, backend: str = "auto", **kw):
    """backend: cpu -> OracleEngine on CPU; cuda -> NativeEngine (fails loudly if the .so is missing)."""
    if backend == "auto":
        backend = "cuda" if torch.cuda.is_available() else "cpu"
    if backend == "cpu":
        kw.pop("precision", None)  # the oracle is fp32
        return OracleEngine(spec, device="cpu", **kw)
    if backend == "cuda":
        from .engine_native import NativeEngine
        return NativeEngine(spec, **kw)
    raise ValueError("unknown backend %r" % backend)
