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
    from bench import BASELINE_IMG_S_PER_REPLICA, ClockSampler
    from tensorflow_distributed_b200.benchmark import _barrier, _l2_flusher, _max_over_ranks, time_steps
    device = torch.device("cuda:%d" % local)
    if world > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=device)
    torch.manual_seed(0)
    model = (CNN() if a.model == "cnn" else MLP(a.hidden, a.layers)).to(device).to(memory_format=torch.channels_last)
    net = nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
    opt = torch.optim.Adam(net.parameters(), lr=0.01)
    B = a.batch
    rng = np.random.RandomState(1234 + rank)
    hx = [torch.from_numpy(rng.rand(B, 784).astype(np.float32)).pin_memory() for _ in range(8)]
    hy = [torch.from_numpy(rng.randint(0, 10, size=B)).pin_memory() for _ in range(8)]
    x, y = hx[0].to(device), hy[0].to(device)

    def step(xx=None, yy=None):
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = F.cross_entropy(net(x if xx is None else xx).float(), y if yy is None else yy)
        loss.backward()
        opt.step()
        return loss

    for _ in range(max(3, a.warmup)):
        step()
    _barrier(world, device)
    flush = None if a.no_l2_flush else _l2_flusher(device)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = _max_over_ranks(time_steps(step, a.steps, flush, device), world, device)
    _barrier(world, device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        l = step(hx[i % 8].to(device, non_blocking=True), hy[i % 8].to(device, non_blocking=True))
        last = float(l)
    e1.record()
    torch.cuda.synchronize(device)
    e2e_ms = _max_over_ranks(e0.elapsed_time(e1), world, device)
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return None
    imgs = world * B * a.steps
    return {"metric": "MNIST-CNN images/sec" if a.model == "cnn" else "MLP images/sec", "impl": "ddp",
            "value": imgs / (ms / 1e3), "unit": "images/s", "n_gpus": world, "steps": a.steps,
            "warmup": max(3, a.warmup), "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": imgs / (ms / 1e3) / BASELINE_IMG_S_PER_REPLICA, "dtype": "bf16 autocast", "data": "synthetic",
            "config": {"model": a.model, "global_batch": world * B, "parallelism": "ddp%d" % world,
                       "stack": "torch eager + cuDNN/cuBLAS + NCCL"},
            "e2e": {"value": imgs / (e2e_ms / 1e3), "unit": "images/s", "h2d_bytes_per_step": B * 784 * 4 + B * 8,
                    "d2h_bytes_per_step": 4, "last_loss": last},
            "gpu_launches": None, "clocks": clocks}
