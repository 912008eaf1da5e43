"""Step tracing (the reference only has time.time() around its loop, m:285-307)."""
import json
import os

from tensorflow_distributed_b200.trainer import run_single
from tensorflow_distributed_b200.utils.trace import StepTracer


def test_disabled_tracer_is_a_no_op():
    t = StepTracer("")
    for s in range(1, 6):
        t.before(s)
        t.after(s)
    t.close()
    assert not t.enabled and t.written is None and t.summary == []


def test_run_single_writes_a_chrome_trace_of_the_requested_window(tmp_path):
    path = str(tmp_path / "steps.json")
    logs = []
    out = run_single(model="mlp", hidden_units=32, training_iters=64 * 9, batch_size=64, display_step=100, nt=1,
                     val_batch=50, backend="cpu", init_stddev=0.1, trace_file=path, trace_start=3, trace_steps=2,
                     log=logs.append)
    assert out["trace"] == path and os.path.exists(path) and os.path.exists(path + ".txt")
    ev = json.load(open(path))["traceEvents"]
    names = {e.get("name", "") for e in ev}
    assert any("addmm" in n or "mm" in n for n in names)  # the oracle's matmuls of the traced steps
    table = open(path + ".txt").read().splitlines()
    assert table[0].split()[0] == "op" and len(table) > 3
    assert any(l.startswith("trace: steps 3..4 written to") for l in logs)


def test_window_that_outlives_the_loop_is_closed_at_the_end(tmp_path):
    path = str(tmp_path / "tail.json")
    out = run_single(model="mlp", hidden_units=16, training_iters=32 * 4, batch_size=32, display_step=100, nt=1,
                     val_batch=50, backend="cpu", init_stddev=0.1, trace_file=path, trace_start=2, trace_steps=50,
                     log=lambda s: None)
    assert out["steps"] == 3 and out["trace"] == path and os.path.getsize(path) > 0
