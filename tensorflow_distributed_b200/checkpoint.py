"""Checkpoint save / restore with the reference's variable names.

The reference has no explicit checkpoint code; saving exists only as a side
effect of ``tf.train.Supervisor(logdir=train_dir)`` writing
``model.ckpt-<global_step>`` from the chief, and ``train_dir`` is a fresh
``mkdtemp`` every launch so nothing is ever resumed
(``/root/reference/mnist_python_m.py:236-253``).  Here it is
explicit: chief-only save, restore-on-start from ``--train_dir``, keys
``global_step``, ``weights/<k>``, ``biases/<k>`` in the reference's layouts
(HWIO filters, [in,out] dense), plus Adam slots, beta powers and RNG state.
``tf_alias_map`` gives TF's auto-names (``Variable`` .. ``Variable_7``) in the
reference's creation order (m:186-195).
"""
from __future__ import annotations

import os
import re
import tempfile
from typing import Dict, Optional

import torch

from .models.spec import ModelSpec

INDEX_FILE = "checkpoint"
PREFIX = "model.ckpt"


def tf_alias_map(spec: ModelSpec) -> Dict[str, str]:
    out = {"global_step": "global_step"}
    for v in spec.variables:
        out[v.name] = v.tf_auto_name
        out[v.name + "/Adam"] = v.tf_auto_name + "/Adam"
        out[v.name + "/Adam_1"] = v.tf_auto_name + "/Adam_1"
    return out


def checkpoint_path(train_dir: str, step: int) -> str:
    return os.path.join(train_dir, "%s-%d.pt" % (PREFIX, step))


def latest_checkpoint(train_dir: str) -> Optional[str]:
    if not train_dir or not os.path.isdir(train_dir):
        return None
    idx = os.path.join(train_dir, INDEX_FILE)
    if os.path.exists(idx):
        with open(idx) as f:
            m = re.search(r'model_checkpoint_path:\s*"([^"]+)"', f.read())
        if m:
            p = os.path.join(train_dir, m.group(1))
            if os.path.exists(p):
                return p
    best, best_step = None, -1
    for fn in os.listdir(train_dir):
        m = re.fullmatch(re.escape(PREFIX) + r"-(\d+)\.pt", fn)
        if m and int(m.group(1)) > best_step:
            best, best_step = os.path.join(train_dir, fn), int(m.group(1))
    return best


def save(train_dir: str, spec: ModelSpec, global_step: int, params: Dict[str, torch.Tensor],
         adam_state: Optional[dict] = None, extra: Optional[dict] = None) -> str:
    os.makedirs(train_dir, exist_ok=True)
    blob = {"format": "tfd_b200.ckpt.v1", "model": spec.kind, "global_step": int(global_step),
            "variables": {}, "tf_alias": tf_alias_map(spec)}
    for v in spec.variables:
        t = params[v.name].detach().to("cpu", torch.float32).contiguous()
        if tuple(t.shape) != tuple(v.shape):
            raise ValueError("variable %s has shape %s, spec says %s" % (v.name, tuple(t.shape), v.shape))
        blob["variables"][v.name] = t
    if adam_state is not None:
        blob["adam"] = {
            "t": int(adam_state["t"]),
            "beta1_power": float(adam_state.get("beta1_power", 0.0)),
            "beta2_power": float(adam_state.get("beta2_power", 0.0)),
            "m": {k: x.detach().to("cpu", torch.float32) for k, x in adam_state["m"].items()},
            "v": {k: x.detach().to("cpu", torch.float32) for k, x in adam_state["v"].items()},
        }
    blob["rng_state"] = torch.get_rng_state()
    if extra:
        blob["extra"] = extra
    path = checkpoint_path(train_dir, global_step)
    fd, tmp = tempfile.mkstemp(dir=train_dir, suffix=".tmp")
    os.close(fd)
    torch.save(blob, tmp)
    os.replace(tmp, path)
    fd, tmp = tempfile.mkstemp(dir=train_dir, suffix=".tmp")
    with os.fdopen(fd, "w") as f:
        f.write('model_checkpoint_path: "%s"\n' % os.path.basename(path))
    os.replace(tmp, os.path.join(train_dir, INDEX_FILE))
    return path


def load(path: str, spec: Optional[ModelSpec] = None) -> dict:
    # weights_only: the blob holds tensors, dicts, strings and numbers only, so nothing in a checkpoint found in
    # train_dir can execute code on load
    blob = torch.load(path, map_location="cpu", weights_only=True)
    if blob.get("format") != "tfd_b200.ckpt.v1":
        raise ValueError("not a tfd_b200 checkpoint: %s" % path)
    if spec is not None:
        for v in spec.variables:
            if v.name not in blob["variables"]:
                raise KeyError("checkpoint %s lacks variable %s" % (path, v.name))
            if tuple(blob["variables"][v.name].shape) != tuple(v.shape):
                raise ValueError("checkpoint variable %s has shape %s, expected %s"
                                 % (v.name, tuple(blob["variables"][v.name].shape), v.shape))
    return blob
