"""Timing helpers.  The reference only has wall-clock ``time.time()`` brackets
(``/root/reference/mnist_python_m.py:285-307``, ``/root/reference/mnist_single.py:102-134``); the trainers keep
those prints, and this module adds device-side timing (CUDA events on the launching stream) plus the reference's
reporting schema ``Steps ,Time ,Accuracy, Learning rate`` (``/root/reference/performance:1``)."""
from __future__ import annotations

import time
from typing import List, Tuple


class DeviceTimer:
    """CUDA-event stopwatch on the current stream; falls back to wall clock without a GPU."""

    def __init__(self):
        import torch
        self._torch = torch
        self._cuda = torch.cuda.is_available()
        self._t0 = None
        if self._cuda:
            self._s, self._e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def start(self) -> None:
        if self._cuda:
            self._s.record()
        else:
            self._t0 = time.perf_counter()

    def stop_ms(self) -> float:
        if self._cuda:
            self._e.record()
            self._e.synchronize()
            return self._s.elapsed_time(self._e)
        return (time.perf_counter() - self._t0) * 1e3


def performance_table(rows: List[Tuple[int, float, float, float]]) -> str:
    """Rows of (steps, seconds, accuracy percent, learning rate) in the layout of the reference's `performance` file."""
    out = ["Steps ,Time ,Accuracy, Learning rate"]
    for steps, secs, acc, lr in rows:
        out.append("%-5d %-9.4g %-6.2f %g" % (steps, secs, acc, lr))
    return "\n".join(out)
