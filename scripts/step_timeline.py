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
from tensorflow_distributed_b200.ops import native as nat  # noqa: E402
adam_s, conv1_s = nat.debug_stamps()  # phase stamps of the last step's Adam / conv1 kernels (this GPU's globaltimer)
if world > 1:
    dist.barrier()
out = os.path.join("gpurun_out", "step_trace_rank%d.json" % rank)
prof.export_chrome_trace(out)
ev = [e for e in json.load(open(out))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy")]
os.remove(out)
ev.sort(key=lambda e: e["ts"])
# last step = kernels after the last fill kernel
idx = max(i for i, e in enumerate(ev) if "fill" in e["name"].lower() or "FillFunctor" in e["name"])
step = ev[idx + 1:]
t0 = step[0]["ts"]
lines = ["%-52s %6s %8s %8s %8s" % ("kernel", "stream", "start", "end", "dur")]
for e in step:
    name = e["name"].replace("(anonymous namespace)::", "").split("(")[0][:52]
    lines.append("%-52s %6s %8.1f %8.1f %8.1f" % (name, e["args"].get("stream", "?"), e["ts"] - t0,
                                                  e["ts"] + e["dur"] - t0, e["dur"]))
lines.append("step span: %.1f us" % (max(e["ts"] + e["dur"] for e in step) - t0))
us = lambda a, b: (b - a) / 1e3
lines.append("conv1 phases (us): gate wait %.1f, compute %.1f" % (us(conv1_s[0], conv1_s[1]) if conv1_s[1] > conv1_s[0] else 0.0,
                                                               us(max(conv1_s[0], conv1_s[1]), conv1_s[2])))
lines.append("adam phases (us): signal %.1f, wait for arrivals %.1f, block-0 update %.1f, all blocks done at %.1f, "
             "release %.1f, total %.1f; conv1 start -> adam start %.1f"
             % (us(adam_s[0], adam_s[1]), us(adam_s[1], adam_s[2]), us(adam_s[2], adam_s[3]), us(adam_s[0], adam_s[4]),
                us(adam_s[4], adam_s[5]) if adam_s[5] > adam_s[4] else 0.0, us(adam_s[0], max(adam_s[4], adam_s[5])),
                us(conv1_s[0], adam_s[0])))
# every rank leaves its own timeline (which rank is late decides how long the others' Adam kernels wait); rank 0 prints
with open(os.path.join("gpurun_out", "timeline_%dgpu_rank%d.txt" % (world, rank)), "w") as f:
    f.write("\n".join(lines) + "\n")
if rank == 0:
    print("\n".join(lines))
if world > 1:
    dist.barrier()
    eng.shutdown()
