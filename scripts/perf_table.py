"""Reproduce the reference's `performance` record (steps / time / accuracy / lr, /root/reference/performance:1-6)
with this framework: train the CNN, stop at 40/60/80/100/120 steps, validate 5x1000 each time.
Synthetic MNIST-shaped data (no network), so accuracies are not comparable to real MNIST; times are."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tensorflow_distributed_b200 import data  # noqa: E402
from tensorflow_distributed_b200.engine import make_engine  # noqa: E402
from tensorflow_distributed_b200.models.spec import cnn_spec  # noqa: E402
from tensorflow_distributed_b200.utils.timing import performance_table  # noqa: E402

backend = sys.argv[1] if len(sys.argv) > 1 else "auto"
ds = data.read_data_sets("", synthetic=True, seed=0)
eng = make_engine(cnn_spec(), backend, learning_rate=0.01, seed=0, init_stddev=1.0)
rows, t0, step = [], time.time(), 0
for stop in (40, 60, 80, 100, 120):
    while step < stop:
        x, y = ds.train.next_batch(128)
        eng.train_step(x, y, 0.75)
        step += 1
    eng.synchronize()
    secs = time.time() - t0
    acc = sum(eng.eval_accuracy(*ds.validation.next_batch(1000)) for _ in range(5)) / 5
    rows.append((stop, secs, 100.0 * acc, 0.01))
print(performance_table(rows))
