"""Supervisor: chief initialises (or restores), non-chiefs wait, chief saves.

What the reference gets from ``tf.train.Supervisor(is_chief, logdir, init_op,
local_init_op, ready_for_local_init_op, recovery_wait_secs=1, global_step)`` and
``sv.prepare_or_wait_for_session`` (``/root/reference/mnist_python_m.py:
236-277``): the chief restores the latest checkpoint in ``logdir`` if there is
one, else runs the initialiser; everybody else polls once per
``recovery_wait_secs`` until the model is ready; a restarted worker re-joins
the same way.  ``train_dir`` defaults to a fresh temp dir as in
the reference (m:236) unless ``--train_dir`` is given, which is what makes
resume possible at all.
"""
from __future__ import annotations

import tempfile
from typing import Callable, Optional

from . import checkpoint
from .models.spec import ModelSpec


class Supervisor:
    def __init__(self, is_chief: bool, spec: ModelSpec, logdir: str = "", recovery_wait_secs: float = 1.0,
                 log: Callable[[str], None] = print):
        self.is_chief = is_chief
        self.spec = spec
        self.logdir = logdir or tempfile.mkdtemp()
        self.recovery_wait_secs = recovery_wait_secs
        self.log = log
        self.restored_from: Optional[str] = None

    def prepare_or_wait(self, *, init_fn: Callable[[], None], restore_fn: Callable[[dict], None],
                        wait_fn: Callable[[float], None]) -> None:
        """Chief: restore-or-init.  Non-chief: wait until the chief has done so."""
        if self.is_chief:
            path = checkpoint.latest_checkpoint(self.logdir)
            if path is not None:
                restore_fn(checkpoint.load(path, self.spec))
                self.restored_from = path
            else:
                init_fn()
        else:
            wait_fn(self.recovery_wait_secs)

    def save(self, global_step: int, params, adam_state=None, extra=None) -> Optional[str]:
        if not self.is_chief:
            return None
        return checkpoint.save(self.logdir, self.spec, global_step, params, adam_state, extra)
