"""Step tracing.  The reference has no profiler hook at all (only ``time.time()`` around the loop,
``/root/reference/mnist_python_m.py:285-307``); here a window of training steps can be recorded and written as a
Chrome/Perfetto trace plus a per-op table:

    python mnist_single.py --trace_file=/tmp/step.json --trace_start=5 --trace_steps=3

On a GPU engine the trace holds the CUDA kernels of the captured step graph (CUPTI keeps per-kernel records for graph
replays; this is what ``profiles/step_timeline.md`` was made from); on the CPU oracle it holds the ATen ops.
Tracing costs time: never read throughput from a traced run.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple


class StepTracer:
    """Record steps ``[start, start + steps)`` (1-based local step numbers).  Call ``before(step)`` / ``after(step)``
    around every training step; everything else is a no-op, and so is the whole object when ``path`` is empty."""

    def __init__(self, path: str = "", start: int = 2, steps: int = 3, log: Callable[[str], None] = print):
        self.path, self.start, self.steps, self.log = path, max(1, int(start)), max(1, int(steps)), log
        self._prof = None
        self.summary: List[Tuple[str, float, int]] = []  # (name, total microseconds, calls)
        self.written: Optional[str] = None

    @property
    def enabled(self) -> bool:
        return bool(self.path)

    def before(self, step: int) -> None:
        if not self.enabled or step != self.start or self._prof is not None or self.written:
            return
        import torch
        from torch.profiler import ProfilerActivity, profile
        acts = [ProfilerActivity.CPU]
        if torch.cuda.is_available():
            torch.cuda.synchronize()
            acts.append(ProfilerActivity.CUDA)
        self._prof = profile(activities=acts)
        self._prof.__enter__()

    def after(self, step: int) -> None:
        if self._prof is None or step < self.start + self.steps - 1:
            return
        self.close()

    def close(self) -> None:
        """Stop recording (also safe to call when the loop ended inside the window) and write the files."""
        if self._prof is None:
            return
        import torch
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        prof, self._prof = self._prof, None
        prof.__exit__(None, None, None)
        prof.export_chrome_trace(self.path)
        cuda = torch.cuda.is_available()
        rows = []
        for e in prof.key_averages():
            us = float(getattr(e, "device_time_total", 0.0) or 0.0) if cuda else float(e.self_cpu_time_total)
            if us > 0:
                rows.append((e.key, us, int(e.count)))
        rows.sort(key=lambda r: -r[1])
        self.summary = rows
        table = self.path + ".txt"
        with open(table, "w") as f:
            f.write("%-60s %12s %8s\n" % ("op / kernel", "total us", "calls"))
            for name, us, n in rows[:40]:
                f.write("%-60s %12.1f %8d\n" % (name[:60], us, n))
        self.written = self.path
        self.log("trace: steps %d..%d written to %s (+ %s)" % (self.start, self.start + self.steps - 1, self.path, table))
