"""Single-process CNN trainer + validation + timing (behaviour of /root/reference/mnist_single.py).

The reference has module constants instead of flags (single:18-26, train_flag single:103); they are
keyword arguments of run_single and can be set as --name=value here.
"""
import sys

from tensorflow_distributed_b200.trainer import run_single

if __name__ == "__main__":
    kw = {}
    casts = {"learning_rate": float, "training_iters": int, "batch_size": int, "display_step": int,
             "dropout": float, "train_flag": int, "nt": int, "val_batch": int, "seed": int,
             "hidden_units": int, "mlp_layers": int, "init_stddev": float, "trace_start": int, "trace_steps": int}
    for a in sys.argv[1:]:
        k, v = a.lstrip("-").split("=", 1)
        if k == "synthetic":
            v = v.lower() in ("1", "true")
        kw[k] = casts.get(k, str)(v) if k != "synthetic" else v
    run_single(**kw)
