"""Single-box launcher: the cluster program with one process per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29500 \
        -m tensorflow_distributed_b200.launch --layout=colocated --train_steps=200

Same flags as the role scripts (``mnist_python_{m,w1,w2}.py``; ``/root/reference/mnist_python_m.py:49-87``) except
that roles come from RANK / WORLD_SIZE and ``--layout`` instead of ``--job_name/--task_index/--*_hosts``:
``colocated`` makes every GPU a worker that also hosts a parameter-server shard, ``dedicated`` makes the first
``--ps_shards`` GPUs ps-only (the reference's topology).  Prints the reference's log lines per worker.
"""
from __future__ import annotations

import os
import sys


def main(argv=None) -> dict:
    import torch
    import torch.distributed as dist

    from . import data as data_mod
    from .config import parse_flags
    from .parallel.ps_gpu import Topology, _gpu_task

    flags = parse_flags(argv, defaults={"job_name": "worker"})
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if not torch.cuda.is_available():
        raise SystemExit("tensorflow_distributed_b200.launch needs GPUs; use the role scripts with host:port lists on CPU")
    layout = "colocated" if flags.layout == "auto" else flags.layout
    topo = Topology.make(world, layout, flags.ps_shards)
    device = torch.device("cuda:%d" % local)
    torch.cuda.set_device(device)
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=device)
    widx = topo.worker_of(rank)
    mnist = data_mod.read_data_sets(flags.data_dir, True, True if flags.synthetic else None,
                                    flags.seed + 101 * (widx if widx is not None else 0))
    print("job name = %s" % ("+".join(n for n, on in (("ps", rank in topo.ps_ranks), ("worker", widx is not None)) if on)))
    print("task index = %d" % (widx if widx is not None else topo.shard_of(rank)))
    out = _gpu_task(flags, topo, rank, device, mnist, print)
    dist.destroy_process_group()
    return out


if __name__ == "__main__":
    main(sys.argv[1:])
