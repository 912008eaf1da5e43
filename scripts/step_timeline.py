"""Kernel timeline of one captured training step (CUPTI through torch.profiler; graph replays keep per-kernel
records).  Prints start/end of every kernel of the last replayed step relative to the step's first kernel."""
import json
import os
import sys

import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tensorflow_distributed_b200.engine_native import NativeEngine  # noqa: E402
from tensorflow_distributed_b200.models.spec import cnn_spec, mlp_spec  # noqa: E402

MODEL = os.environ.get("TFD_MODEL", "cnn")
SPEC = cnn_spec() if MODEL == "cnn" else mlp_spec(int(os.environ.get("TFD_HIDDEN", "8192")), int(os.environ.get("TFD_LAYERS", "4")))
KEEP = 0.75 if MODEL == "cnn" else 1.0
STD = 1.0 if MODEL == "cnn" else 0.05

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
if world > 1:  # under torchrun: colocated ps shards, one rank per GPU; rank 0 prints its own timeline
    import torch.distributed as dist
    from tensorflow_distributed_b200.parallel.ps_gpu import GpuClusterEngine, Topology
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % local))
    eng = GpuClusterEngine(SPEC, Topology.make(world, os.environ.get("TFD_LAYOUT", "colocated")), rank,
                           "cuda:%d" % local, init_stddev=STD, pull_mode=os.environ.get("TFD_PULL", "auto"),
                           precision=os.environ.get("TFD_PRECISION", "tf32"))
else:
    eng = NativeEngine(SPEC, init_stddev=STD, precision=os.environ.get("TFD_PRECISION", "tf32"))
rng = np.random.RandomState(rank)
x = rng.rand(128, 784).astype(np.float32)
y = np.eye(10, dtype=np.float32)[rng.randint(0, 10, size=128)]
ws = eng.workspace(128)
eng.load_batch(ws, x, y)
for _ in range(5):
    eng.train_step_resident(ws, KEEP)
torch.cuda.synchronize()
flush = torch.empty(64 * 1024 * 1024, device="cuda")
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        flush.fill_(1.0)
        eng.train_step_resident(ws, KEEP)
    torch.cuda.synchronize()
if world > 1:
    dist.barrier()
if rank != 0:
    eng.shutdown()
    sys.exit(0)
out = os.path.join("gpurun_out", "step_trace.json")
prof.export_chrome_trace(out)
ev = [e for e in json.load(open(out))["traceEvents"] if e.get("cat") == "kernel"]
ev.sort(key=lambda e: e["ts"])
# last step = kernels after the last fill kernel
idx = max(i for i, e in enumerate(ev) if "fill" in e["name"].lower() or "FillFunctor" in e["name"])
step = ev[idx + 1:]
t0 = step[0]["ts"]
print("%-52s %6s %8s %8s %8s" % ("kernel", "stream", "start", "end", "dur"))
for e in step:
    name = e["name"].replace("(anonymous namespace)::", "").split("(")[0][:52]
    print("%-52s %6s %8.1f %8.1f %8.1f" % (name, e["args"].get("stream", "?"), e["ts"] - t0, e["ts"] + e["dur"] - t0, e["dur"]))
print("step span: %.1f us" % (step[-1]["ts"] + step[-1]["dur"] - t0))
if world > 1:
    eng.shutdown()
