"""Replica update protocols: synchronous (aggregate-R-then-apply) and asynchronous
(apply-on-arrival).

Pure host-side state machines with no transport in them.  They are (a) the
logic the CPU/gloo parameter server runs, (b) the executable specification the
device-side protocol (epoch flags + arrival counters in peer-visible memory,
``csrc/ps_kernels.cu``) is tested against, and (c) the model used by the unit
tests of SURVEY.md section 4.

Semantics follow what the reference configures through
``tf.train.SyncReplicasOptimizer`` (``/root/reference/mnist_python_m.py:210-233,
279-282``) and plain ``AdamOptimizer.minimize`` for the async branch (m:208, 222):

* one gradient accumulator per variable on the ps; a gradient tagged with a
  ``local_step`` older than ``global_step`` is dropped (stale);
* when every accumulator holds R gradients the chief's aggregation takes their
  *average*, applies the optimizer once, increments ``global_step`` and releases
  N tokens carrying the new global step; each worker dequeues one token per
  step and adopts it as its ``local_step``;
* ``get_init_tokens_op()`` preloads R tokens (m:233, 281) so workers may run one
  step ahead of the first aggregation;
* async: every arriving gradient is applied at once, in a definite per-shard
  order (ticket), and every apply increments ``global_step`` (m:301 counts
  worker updates in that mode).
"""
from __future__ import annotations

from collections import deque
from typing import Callable, Deque, Dict, List, Optional

import torch


class ConditionalAccumulator:
    """Gradient accumulator with stale-drop (the TF op of the same name)."""

    def __init__(self, name: str):
        self.name = name
        self.sum: Optional[torch.Tensor] = None
        self.count = 0
        self.dropped = 0
        self.global_step = 0

    def set_global_step(self, step: int) -> None:
        self.global_step = step

    def apply_grad(self, grad: torch.Tensor, local_step: int) -> bool:
        if local_step < self.global_step:
            self.dropped += 1
            return False
        self.sum = grad.clone() if self.sum is None else self.sum.add_(grad)
        self.count += 1
        return True

    def ready(self, required: int) -> bool:
        return self.count >= required

    def take_grad(self, required: int) -> torch.Tensor:
        """Average of everything accumulated (>= required); resets the accumulator."""
        assert self.ready(required), "take_grad before %d gradients arrived" % required
        avg = self.sum / float(self.count)
        self.sum, self.count = None, 0
        return avg


class SyncReplicas:
    """Aggregate ``replicas_to_aggregate`` gradients, apply once, release ``total_num_replicas`` tokens."""

    def __init__(self, var_names: List[str], replicas_to_aggregate: int, total_num_replicas: int,
                 apply_fn: Callable[[Dict[str, torch.Tensor]], None], global_step: int = 0):
        if replicas_to_aggregate < 1 or total_num_replicas < 1:
            raise ValueError("replica counts must be >= 1")
        self.R = replicas_to_aggregate
        self.N = total_num_replicas
        self.apply_fn = apply_fn
        self.global_step = global_step
        self.acc = {n: ConditionalAccumulator(n) for n in var_names}
        for a in self.acc.values():
            a.set_global_step(global_step)
        self.tokens: Deque[int] = deque()
        self.applied = 0

    # -- chief side -------------------------------------------------------
    def init_tokens(self, num_tokens: int = -1) -> None:
        """get_init_tokens_op(): default preload is R tokens of the current global step."""
        n = self.R if num_tokens < 0 else num_tokens
        if self.N > self.R and n < self.N - self.R:
            raise ValueError("too few init tokens: workers beyond R would starve")
        for _ in range(n):
            self.tokens.append(self.global_step)

    def try_aggregate(self) -> bool:
        """One turn of the chief queue runner; True if an update was applied."""
        if not all(a.ready(self.R) for a in self.acc.values()):
            return False
        grads = {n: a.take_grad(self.R) for n, a in self.acc.items()}
        self.apply_fn(grads)
        self.applied += 1
        self.global_step += 1
        for a in self.acc.values():
            a.set_global_step(self.global_step)
        for _ in range(self.N):
            self.tokens.append(self.global_step)
        return True

    # -- worker side ------------------------------------------------------
    def push(self, grads: Dict[str, torch.Tensor], local_step: int) -> int:
        """Returns how many of the gradients were accepted (0 if the whole push was stale)."""
        accepted = 0
        for n, g in grads.items():
            accepted += bool(self.acc[n].apply_grad(g, local_step))
        return accepted

    def token_available(self) -> bool:
        return len(self.tokens) > 0

    def dequeue_token(self) -> int:
        return self.tokens.popleft()

    @property
    def dropped(self) -> int:
        return sum(a.dropped for a in self.acc.values())


class AsyncReplicas:
    """Apply-on-arrival with a per-shard ticket: every gradient applied exactly once, in ticket order."""

    def __init__(self, apply_fn: Callable[[Dict[str, torch.Tensor]], None], global_step: int = 0):
        self.apply_fn = apply_fn
        self.global_step = global_step
        self.next_ticket = 0
        self.served = 0
        self.order: List[int] = []  # worker id per applied ticket

    def take_ticket(self) -> int:
        t = self.next_ticket
        self.next_ticket += 1
        return t

    def push(self, worker: int, grads: Dict[str, torch.Tensor], ticket: Optional[int] = None) -> int:
        if ticket is None:
            ticket = self.take_ticket()
        if ticket != self.served:
            raise RuntimeError("ticket %d applied out of order (serving %d)" % (ticket, self.served))
        self.apply_fn(grads)
        self.served += 1
        self.global_step += 1
        self.order.append(worker)
        return self.global_step
