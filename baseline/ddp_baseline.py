"""Second baseline (BASELINE.json north star): the same model in stock PyTorch -- cuDNN/cuBLAS kernels, autograd,
torch.optim.Adam, DistributedDataParallel over NCCL when N > 1 -- timed with the same harness as bench.py.
None of this repo's kernels, engine or ps protocol is on this path."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


class CNN(nn.Module):
    def __init__(self):
        super().__init__()
        self.c1 = nn.Conv2d(1, 32, 5, padding=2)
        self.c2 = nn.Conv2d(32, 64, 5, padding=2)
        self.f1 = nn.Linear(3136, 1024)
        self.out = nn.Linear(1024, 10)
        for p in self.parameters():
            nn.init.normal_(p, 0.0, 1.0)

    def forward(self, x):
        x = x.view(-1, 1, 28, 28)
        x = F.max_pool2d(F.relu(self.c1(x)), 2)
        x = F.max_pool2d(F.relu(self.c2(x)), 2)
        x = x.permute(0, 2, 3, 1).reshape(-1, 3136)
        x = F.dropout(F.relu(self.f1(x)), 0.25, self.training)
        return self.out(x)


class MLP(nn.Module):
    def __init__(self, hidden, layers):
        super().__init__()
        dims = [784] + [hidden] * layers
        self.hid = nn.ModuleList(nn.Linear(dims[i], dims[i + 1]) for i in range(layers))
        self.out = nn.Linear(dims[-1], 10)

    def forward(self, x):
        for l in self.hid:
            x = F.relu(l(x))
        return self.out(x)


def run_ddp(a, rank, world, local):
    """--impl ddp [--graphs] [--precision tf32|bf16] [--compile].

    Eager arm: torch DistributedDataParallel (bucketed NCCL all-reduce overlapped with backward) + torch.optim.Adam.
    --graphs arm: the whole step -- forward, backward, ONE flat-bucket NCCL all-reduce of the gradients (averaged),
    Adam(capturable) -- is captured in a CUDA graph with static inputs and replayed, which is the like-for-like
    baseline for this repo's graph-replayed step (an eager 10-GFLOP step is pure launch overhead).
    precision tf32: fp32 tensors, TF32 tensor cores allowed for conv / matmul; bf16: autocast."""
    from bench import ClockSampler
    from tensorflow_distributed_b200.benchmark import (RankAligner, _barrier, _l2_flusher, _max_over_ranks, _stats,
                                                       time_rounds, time_steps)
    import torch.distributed as dist
    device = torch.device("cuda:%d" % local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=device)
    precision = a.precision or "tf32"
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(0)
    model = (CNN() if a.model == "cnn" else MLP(a.hidden, a.layers)).to(device).to(memory_format=torch.channels_last)
    fwd = torch.compile(model) if getattr(a, "compile", False) else model
    use_graphs = bool(getattr(a, "graphs", False))
    net = fwd
    if world > 1 and not use_graphs:
        net = nn.parallel.DistributedDataParallel(fwd, device_ids=[local])
    params = [p for p in model.parameters()]
    opt = torch.optim.Adam(params, lr=0.01, capturable=use_graphs)
    B = a.batch
    rng = np.random.RandomState(1234 + rank)
    hx = [torch.from_numpy(rng.rand(B, 784).astype(np.float32)).pin_memory() for _ in range(8)]
    hy = [torch.from_numpy(rng.randint(0, 10, size=B)).pin_memory() for _ in range(8)]
    x, y = hx[0].to(device), hy[0].to(device)
    loss_out = torch.zeros((), device=device)

    def body():
        opt.zero_grad(set_to_none=False)
        if precision == "bf16":
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = F.cross_entropy(net(x).float(), y)
        else:
            loss = F.cross_entropy(net(x), y)
        loss.backward()
        if use_graphs and world > 1:  # DDP's all-reduce as one flat bucket inside the captured step
            flat = torch._utils._flatten_dense_tensors([p.grad for p in params])
            dist.all_reduce(flat)
            flat.div_(world)
            for p, g in zip(params, torch._utils._unflatten_dense_tensors(flat, [p.grad for p in params])):
                p.grad.copy_(g)
        opt.step()
        loss_out.copy_(loss.detach())

    if use_graphs:
        side = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(11):
                body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize(device)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            body()
        step = graph.replay
    else:
        step = body

    def step_host(i):
        x.copy_(hx[i % 8], non_blocking=True)
        y.copy_(hy[i % 8], non_blocking=True)
        step()
        return float(loss_out)  # D2H of the loss every step

    for _ in range(max(3, a.warmup)):
        step()
    _barrier(world, device)
    flush = None if a.no_l2_flush else _l2_flusher(device)
    align = RankAligner(world, device)
    time_steps(step, 3, flush, device, align)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    sums, per_step = time_rounds(step, a.steps, flush, device, align, world, a.min_seconds)
    ms = sum(sums) / len(sums)
    e2e_rounds = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    last = 0.0
    for _ in range(max(1, min(len(sums), 20))):
        _barrier(world, device)
        e0.record()
        for i in range(a.steps):
            last = step_host(i)
        e1.record()
        torch.cuda.synchronize(device)
        e2e_rounds.append(_max_over_ranks(e0.elapsed_time(e1), world, device))
    e2e_ms = sum(e2e_rounds) / len(e2e_rounds)
    clocks = sampler.stop() if rank == 0 else None
    _barrier(world, device)
    if rank != 0:
        return None
    imgs = world * B * a.steps
    stack = ("CUDA-graph replay of fwd+bwd+flat NCCL all-reduce+Adam(capturable)" if use_graphs
             else "torch eager + DistributedDataParallel + torch.optim.Adam")
    return {"metric": "MNIST-CNN images/sec" if a.model == "cnn" else "MLP images/sec",
            "impl": "ddp-graph" if use_graphs else "ddp",
            "value": imgs / (ms / 1e3), "unit": "images/s", "n_gpus": world, "steps": a.steps,
            "warmup": max(3, a.warmup), "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "tf32 (fp32 tensors, allow_tf32)" if precision == "tf32" else "bf16 autocast",
            "data": "synthetic",
            "config": {"model": a.model, "global_batch": world * B, "parallelism": "ddp%d" % world,
                       "stack": stack + "; cuDNN/cuBLAS/NCCL" + ("; torch.compile" if getattr(a, "compile", False) else ""),
                       "rounds": len(sums), "round_ms": _stats(sums)},
            "e2e": {"value": imgs / (e2e_ms / 1e3), "unit": "images/s", "h2d_bytes_per_step": B * 784 * 4 + B * 8,
                    "d2h_bytes_per_step": 4, "last_loss": last, "ms_per_step": e2e_ms / a.steps},
            "gpu_launches": None, "clocks": clocks}
