"""Model variable specs (names, shapes, creation order).

CNN: the reference's 2-conv + 2-FC network, parameter dicts ``weights{wc1,wc2,
wd1,out}`` / ``biases{bc1,bc2,bd1,out}`` (``/root/reference/mnist_python_m.py:
185-196``), created in that order after ``global_step`` (m:178).  Canonical
checkpoint keys are ``weights/<k>`` and ``biases/<k>`` in the reference's
layouts (HWIO filters, [in,out] dense, NHWC flatten order).

MLP: not in the reference source (only the docstring m:18 and the dead flag
``hidden_units`` m:66-67 survive).  784 -> hidden_units (x mlp_layers) -> 10,
named in the reference's dict style: ``weights/h1..hN``, ``weights/out``,
``biases/b1..bN``, ``biases/out``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Tuple

IMAGE_PIXELS = 28  # m:90
N_INPUT = IMAGE_PIXELS * IMAGE_PIXELS
N_CLASSES = 10


@dataclass(frozen=True)
class VariableSpec:
    name: str                 # canonical checkpoint key
    shape: Tuple[int, ...]
    tf_auto_name: str = ""    # what TF 1.x would have auto-named it (Variable, Variable_1, ...)

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n

    @property
    def rows(self) -> int:
        """Leading extent used for intra-tensor (row) sharding: product of all but the last dim."""
        n = 1
        for s in self.shape[:-1]:
            n *= s
        return n if len(self.shape) > 1 else self.shape[0]

    @property
    def cols(self) -> int:
        return self.shape[-1] if len(self.shape) > 1 else 1


@dataclass(frozen=True)
class ModelSpec:
    kind: str
    variables: Tuple[VariableSpec, ...]
    hidden: Tuple[int, ...] = ()

    @property
    def num_params(self) -> int:
        return sum(v.numel for v in self.variables)

    def by_name(self) -> Dict[str, VariableSpec]:
        return {v.name: v for v in self.variables}

    def names(self) -> List[str]:
        return [v.name for v in self.variables]


def _tf_name(i: int) -> str:
    return "Variable" if i == 0 else "Variable_%d" % i


def cnn_spec() -> ModelSpec:
    shapes = [
        ("weights/wc1", (5, 5, 1, 32)),
        ("weights/wc2", (5, 5, 32, 64)),
        ("weights/wd1", (7 * 7 * 64, 1024)),
        ("weights/out", (1024, N_CLASSES)),
        ("biases/bc1", (32,)),
        ("biases/bc2", (64,)),
        ("biases/bd1", (1024,)),
        ("biases/out", (N_CLASSES,)),
    ]
    return ModelSpec("cnn", tuple(VariableSpec(n, s, _tf_name(i)) for i, (n, s) in enumerate(shapes)))


def mlp_spec(hidden_units: int = 100, layers: int = 1) -> ModelSpec:
    dims = [N_INPUT] + [hidden_units] * layers
    ws, bs = [], []
    for i in range(layers):
        ws.append(("weights/h%d" % (i + 1), (dims[i], dims[i + 1])))
        bs.append(("biases/b%d" % (i + 1), (dims[i + 1],)))
    ws.append(("weights/out", (dims[-1], N_CLASSES)))
    bs.append(("biases/out", (N_CLASSES,)))
    allv = ws + bs
    return ModelSpec("mlp", tuple(VariableSpec(n, s, _tf_name(i)) for i, (n, s) in enumerate(allv)),
                     hidden=tuple(dims[1:]))


def make_spec(kind: str, hidden_units: int = 100, layers: int = 1) -> ModelSpec:
    if kind == "cnn":
        return cnn_spec()
    if kind == "mlp":
        return mlp_spec(hidden_units, layers)
    raise ValueError("unknown model %r" % kind)
