"""A/B check of the experimental NVLS transport against the default peer-memory transport, under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
        scripts/nvls_check.py [--layout dedicated] [--model mlp]

Both engines start from the same seed and see the same batches; after a few synchronous steps their parameters
and Adam slots must agree (the only differences are the order of the floating-point gradient sum -- in the
switch instead of in the shard's Adam kernel -- and atomics in the conv weight-gradient kernels)."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tensorflow_distributed_b200 import data  # noqa: E402
from tensorflow_distributed_b200.models.spec import make_spec  # noqa: E402
from tensorflow_distributed_b200.parallel.ps_gpu import GpuClusterEngine, Topology  # noqa: E402
from tensorflow_distributed_b200.parallel.ps_nvls import NvlsClusterEngine  # noqa: E402


def rel(a, b):
    return (a.float() - b.float()).norm().item() / (b.float().norm().item() + 1e-20)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layout", default="colocated")
    ap.add_argument("--ps", type=int, default=0)
    ap.add_argument("--model", default="cnn")
    ap.add_argument("--hidden", type=int, default=512)
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--precision", default=None, choices=["tf32", "bf16"])
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    device = torch.device("cuda:%d" % local)
    dist.init_process_group("nccl", device_id=device)
    spec = make_spec(a.model, a.hidden, a.layers)
    topo = Topology.make(world, a.layout, a.ps)
    w = topo.worker_of(rank)
    stream = data.read_data_sets("", synthetic=True, seed=100 + (w if w is not None else 99)).train
    batches = [stream.next_batch(128) for _ in range(a.steps)]
    results = []
    for cls in (GpuClusterEngine, NvlsClusterEngine):
        eng = cls(spec, topo, rank, device, learning_rate=0.01, seed=3, init_stddev=0.05, precision=a.precision)
        losses = [eng.train_step(x, y, 1.0, fetch_loss=True) for x, y in batches]  # eager, eager, then graph replays
        torch.cuda.synchronize()
        dist.barrier()
        results.append((eng.get_params(), eng.get_adam_state(), eng.global_step, losses, eng.describe()))
        eng.shutdown()
    (p0, a0, g0, l0, d0), (p1, a1, g1, l1, d1) = results
    ok, msgs = True, []
    if not (g0 == g1 == a.steps):
        ok = False
        msgs.append("global steps %d / %d, expected %d" % (g0, g1, a.steps))
    for k in p0:
        for what, x, y in (("param", p1[k], p0[k]), ("m", a1["m"][k], a0["m"][k]), ("v", a1["v"][k], a0["v"][k])):
            e = rel(x, y)
            if e > 2e-3:
                ok = False
                msgs.append("%s %s rel err %.2e" % (what, k, e))
    if w is not None:  # the same trajectory on both transports (the gradient sum order is the only difference)
        for i, (x, y) in enumerate(zip(l1, l0)):
            if abs(x - y) > 2e-3 * max(1.0, abs(y)):
                ok = False
                msgs.append("step %d loss %.5f vs %.5f on the peer transport" % (i, x, y))
                break
    flag = torch.tensor([0 if ok else 1], device=device)
    dist.all_reduce(flag)
    print("rank %d [%s] vs [%s] %s losses %.4f -> %.4f (peer %.4f -> %.4f) %s" % (
        rank, d1, d0, "OK" if ok else "FAIL", l1[0] if l1 and l1[0] is not None else -1, l1[-1] if l1 and l1[-1] is not None else -1,
        l0[0] if l0 and l0[0] is not None else -1, l0[-1] if l0 and l0[-1] is not None else -1, "; ".join(msgs)), flush=True)
    dist.destroy_process_group()
    sys.exit(1 if flag.item() else 0)


if __name__ == "__main__":
    main()
