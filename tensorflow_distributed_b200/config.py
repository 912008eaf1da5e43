"""Flag / config system.

Mirrors the reference's 13 ``tf.app.flags`` (same names, same defaults:
``/root/reference/mnist_python_m.py:49-87``) and adds the selectors this
framework needs (model, ps sharding, train_dir, synthetic data, precision).

The reference ships three copies of one program that differ only in the
defaults of ``job_name`` and ``task_index`` (``mnist_python_m.py:55,85``,
``mnist_python_w1.py:85``, ``mnist_python_w2.py:55``).  ``role_defaults``
reproduces those three default sets so the three thin launcher scripts at the
repo root can share one implementation.

TF-style boolean syntax is accepted: ``--sync_replicas=False``,
``--sync_replicas false``, ``--nosync_replicas`` and bare ``--sync_replicas``.
"""
from __future__ import annotations

import dataclasses
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence


@dataclass
class Flags:
    # ---- the reference's 13 flags (mnist_python_m.py:49-87), same defaults ----
    data_dir: str = "/tmp/mnist-data"
    download_only: bool = False
    task_index: int = 0
    num_gpus: int = 0
    replicas_to_aggregate: Optional[int] = 2
    hidden_units: int = 100
    train_steps: int = 4
    batch_size: int = 128
    learning_rate: float = 0.01
    sync_replicas: bool = True
    existing_servers: bool = False
    ps_hosts: str = "10.0.1.3:2222"
    worker_hosts: str = "10.0.1.6:2223,10.0.1.2:2224"
    job_name: str = "ps"
    # ---- additions of this framework ----
    model: str = "cnn"              # cnn | mlp
    mlp_layers: int = 1             # hidden layers of the MLP (4 for the scaled benchmark)
    ps_shards: int = 0              # 0 => len(ps_hosts)
    synthetic: bool = True          # no network on the box: synthetic MNIST-shape data by default
    train_dir: str = ""             # "" => fresh tempfile.mkdtemp() like the reference (m:236)
    precision: str = "tf32"         # tensor-core operand type on the GPU path: tf32 (fp32 in memory) | bf16
    seed: int = 0
    keep_prob: float = 0.75         # m:292
    backend: str = "auto"           # auto | cpu | cuda
    save_checkpoint_steps: int = 0  # 0 => only at end of training
    validation_batches: int = 5     # m:311
    validation_batch_size: int = 1000  # m:313
    init_stddev: float = 1.0        # tf.random_normal default (m:186-195)
    step_timeout_secs: float = 120.0   # watchdog on blocking waits
    inject_fault: str = ""          # "", "stall:<task>:<secs>", "kill:<task>:<step>" (tests)
    log_device_placement: bool = False  # m:257
    sync_init_tokens: int = -1      # get_init_tokens_op(num_tokens); -1 => replicas_to_aggregate (m:233)
    layout: str = "auto"            # GPU cluster: auto | colocated | dedicated ps GPUs
    grad_dtype: str = "fp32"        # dtype of pushed gradients on the GPU path: fp32 | bf16
    transport: str = "auto"         # GPU cluster transport: auto | peer (cudaIpc peer memory) | nvls (NVSwitch multicast)
    trace_file: str = ""            # write a Chrome trace of a window of steps (utils/trace.py)
    trace_start: int = 2            # first local step of the window (1-based)
    trace_steps: int = 3            # length of the window
    server_generations: int = 1     # ps task: how many successive sets of clients it serves before it exits (0: until
                                    # a SHUTDOWN message); later sets attach with --existing_servers=True

    def replace(self, **kw) -> "Flags":
        return dataclasses.replace(self, **kw)


#: Per-script defaults of the three reference launchers.
ROLE_DEFAULTS: Dict[str, Dict[str, Any]] = {
    "m": {"job_name": "ps", "task_index": 0},       # mnist_python_m.py:55,85
    "w1": {"job_name": "worker", "task_index": 0},  # mnist_python_w1.py:55,85
    "w2": {"job_name": "worker", "task_index": 1},  # mnist_python_w2.py:55,85
}

_HELP = {
    "data_dir": "Directory for storing mnist data",
    "download_only": "Only load/prepare the data; do not build the model or train",
    "task_index": "Worker task index, >= 0. task_index=0 is the chief that initialises variables",
    "num_gpus": "Total number of gpus for each machine. 0 => CPU",
    "replicas_to_aggregate": "Replicas to aggregate before a parameter update is applied "
                             "(sync_replicas mode only; 0/None => num_workers)",
    "hidden_units": "Number of units in the hidden layer of the MLP",
    "train_steps": "Number of (global) training steps to perform",
    "batch_size": "Training batch size (per worker)",
    "learning_rate": "Learning rate",
    "sync_replicas": "Aggregate worker updates before applying (sync) instead of apply-on-arrival (async)",
    "existing_servers": "Attach to already-running ps servers instead of starting an in-process one",
    "ps_hosts": "Comma-separated list of host:port (CPU/gloo) or gpu:<ordinal> (single B200 box)",
    "worker_hosts": "Comma-separated list of host:port or gpu:<ordinal>",
    "job_name": "job name: worker or ps",
}


#: Help for the flags this framework adds (kept apart: ``_HELP`` is exactly the reference's flag set).
_HELP_EXTRA = {
    "model": "cnn (the reference's conv_net) or mlp (784 -> hidden_units x mlp_layers -> 10)",
    "mlp_layers": "Hidden layers of the MLP",
    "ps_shards": "Parameter shards (0 => one per ps task)",
    "synthetic": "Use the deterministic MNIST-shaped synthetic set instead of idx files in data_dir",
    "train_dir": "Checkpoint directory (chief saves, everybody restores on start); '' => fresh temp dir",
    "precision": "Tensor-core operand type of the sm_100a kernels: tf32 (fp32 activations/weights read as tf32; the "
                 "arm comparable to the reference's fp32 graph) | bf16 (bf16 activations + bf16 weight shadow); master "
                 "weights, Adam slots, gradients and accumulation are fp32 in both",
    "seed": "Seed of the parameter init, the data stream and dropout",
    "keep_prob": "Dropout keep probability during training",
    "backend": "auto | cpu (fp32 oracle) | cuda (sm_100a kernels)",
    "save_checkpoint_steps": "Chief also saves every N global steps (0 => only when training ends)",
    "validation_batches": "Validation batches after training",
    "validation_batch_size": "Images per validation batch",
    "init_stddev": "Standard deviation of the random_normal initialiser",
    "step_timeout_secs": "Watchdog on every blocking wait of the cluster protocol",
    "inject_fault": "Testing: stall:<task>:<secs> or kill:<task>:<local step>",
    "log_device_placement": "Print the variable -> ps shard placement table",
    "sync_init_tokens": "Tokens preloaded into the sync queue (-1 => replicas_to_aggregate)",
    "layout": "GPU cluster: colocated (every GPU is worker + ps shard) or dedicated ps GPUs",
    "transport": "GPU cluster transport: peer (cudaIpc peer memory, unicast), nvls (NVSwitch multicast: gradient sum in the "
                 "switch, one-store parameter publish; strict sync only) or auto (nvls on dedicated ps GPUs when the box "
                 "has multicast, else peer)",
    "server_generations": "ps task (host:port clusters): serve this many successive sets of worker processes, keeping "
                          "variables, Adam slots and global_step between them (0 = until a SHUTDOWN message); the later "
                          "sets attach with --existing_servers=True (m:76-80, m:268-273)",
    "grad_dtype": "Storage type of pushed gradients in the ps staging slots on the GPU path: fp32 | bf16",
    "trace_file": "Write a Chrome trace (+ .txt per-op table) of a window of training steps",
    "trace_start": "First local step of the traced window (1-based)",
    "trace_steps": "Number of traced steps",
}


def _to_bool(s: str) -> bool:
    v = s.strip().lower()
    if v in ("1", "true", "t", "yes", "y"):
        return True
    if v in ("0", "false", "f", "no", "n"):
        return False
    raise ValueError("not a boolean: %r" % s)


def flag_names() -> List[str]:
    return [f.name for f in dataclasses.fields(Flags)]


def reference_flag_names() -> List[str]:
    """The 13 flags of the reference plus job_name (14 definitions, m:49-85)."""
    return list(_HELP.keys())


def help_text() -> str:
    lines = ["flags:"]
    d = Flags()
    for f in dataclasses.fields(Flags):
        lines.append("  --%s (default %r) %s" % (f.name, getattr(d, f.name), _HELP.get(f.name) or _HELP_EXTRA.get(f.name, "")))
    return "\n".join(lines)


def parse_flags(argv: Optional[Sequence[str]] = None, defaults: Optional[Dict[str, Any]] = None) -> Flags:
    """Parse ``--name=value`` / ``--name value`` / ``--[no]bool`` into Flags."""
    import sys

    argv = list(sys.argv[1:] if argv is None else argv)
    flags = Flags(**(defaults or {}))
    fields = {f.name: f for f in dataclasses.fields(Flags)}
    i = 0
    while i < len(argv):
        tok = argv[i]
        i += 1
        if tok in ("-h", "--help"):
            print(help_text())
            raise SystemExit(0)
        if not tok.startswith("--"):
            raise SystemExit("unexpected argument %r\n%s" % (tok, help_text()))
        body = tok[2:]
        if "=" in body:
            name, val = body.split("=", 1)
        else:
            name, val = body, None
        if name not in fields and name.startswith("no") and name[2:] in fields \
                and isinstance(getattr(Flags(), name[2:]), bool):
            setattr(flags, name[2:], False)
            continue
        if name not in fields:
            raise SystemExit("unknown flag --%s\n%s" % (name, help_text()))
        cur = getattr(Flags(), name)
        if isinstance(cur, bool):
            if val is None:
                # bare --flag, or "--flag false" as a separate token
                if i < len(argv) and not argv[i].startswith("--"):
                    try:
                        b = _to_bool(argv[i])
                        i += 1
                        setattr(flags, name, b)
                        continue
                    except ValueError:
                        pass
                setattr(flags, name, True)
            else:
                setattr(flags, name, _to_bool(val))
            continue
        if val is None:
            if i >= len(argv):
                raise SystemExit("flag --%s needs a value" % name)
            val = argv[i]
            i += 1
        if name == "replicas_to_aggregate":
            setattr(flags, name, None if val in ("", "None", "none") else int(val))
        elif isinstance(cur, int):
            setattr(flags, name, int(val))
        elif isinstance(cur, float):
            setattr(flags, name, float(val))
        else:
            setattr(flags, name, val)
    return flags
