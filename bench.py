"""Headline benchmark: MNIST-CNN training images/sec on N B200s (BASELINE.json metric).

    python bench.py --gpus 1 --steps 50 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

One JSON line on rank 0.  `value` is device-timed (CUDA events per step, L2 flushed between steps, ranks aligned
by an untimed device-side rendezvous in front of every step, max over ranks; the K-step measurement is repeated
for >= --min-seconds and the mean round is reported); `e2e` goes through the public API (`engine.train_step(numpy_x, numpy_y, fetch_loss=True)`): pinned-host
H2D of the inputs and a D2H of the loss every step.  Synthetic 28x28 data, random-init weights (reference init
N(0,1), /root/reference/mnist_python_m.py:186-195), batch 128 per worker (m:70), Adam lr 0.01 (m:71), keep 0.75.

--impl reference: the reference is six TensorFlow-1.x scripts with nothing installable and TensorFlow is not in
this image (no network), so that arm reports itself unavailable.  --impl ddp is this repo's own second baseline:
the same CNN in stock PyTorch (cuDNN/cuBLAS, autograd, torch.optim.Adam, DDP over NCCL when N > 1).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)



class ClockSampler(threading.Thread):
    """SM clock and throttle reasons sampled DURING the timed regions (B200_PROFILING.md recipe): NVML every 5 ms
    (the same counters nvidia-smi prints; spawning nvidia-smi takes longer than a whole timed region)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.sm, self.max_sm, self.reasons, self._halt = index, [], None, set(), threading.Event()
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            uuid = None
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
        try:
            r = n.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20),
                          ("hw_thermal_slowdown", 0x40)):
            if r & bit:
                self.reasons.add(name)

    def _sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout
        c = [x.strip() for x in out.strip().split(",")]
        self.sm.append(float(c[0])); self.max_sm = float(c[1])
        for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], c[2:6]):
            if v.lower().startswith("active"):
                self.reasons.add(name)

    def run(self):
        while not self._halt.is_set():
            try:
                self._sample_nvml() if self.nvml else self._sample_smi()
            except Exception:
                pass
            self._halt.wait(0.002 if self.nvml else 0.2)

    def stop(self) -> dict:
        self._halt.set()
        self.join(timeout=6)
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_sm, "reasons": sorted(self.reasons),
                "samples": len(sm), "source": "nvml" if self.nvml else "nvidia-smi"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "ddp"])
    ap.add_argument("--model", default="cnn", choices=["cnn", "mlp"])
    ap.add_argument("--hidden", type=int, default=8192)
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--layout", default="colocated", choices=["colocated", "dedicated"])
    ap.add_argument("--ps", type=int, default=0, help="ps shards (0: one per GPU when colocated, 1 when dedicated)")
    ap.add_argument("--async", dest="async_mode", action="store_true")
    ap.add_argument("--pull", default="auto", choices=["auto", "fused", "prefetch"],
                    help="big-matrix pull: fused into the GEMMs (TMA from the ps GPU) or prefetched under the conv layers")
    ap.add_argument("--grad-dtype", dest="grad_dtype", default="fp32", choices=["fp32", "bf16"],
                    help="storage type of the pushed gradients in the ps staging slots (multi-GPU)")
    ap.add_argument("--transport", default=None, choices=["auto", "peer", "nvls"],
                    help="multi-GPU ps transport: peer (cudaIpc peer memory, unicast) or nvls (NVSwitch multicast: gradient "
                         "sum in the switch, one-store parameter publish); default: env TFD_TRANSPORT or auto (nvls on dedicated "
                         "ps GPUs when the box has multicast)")
    ap.add_argument("--no-l2-flush", action="store_true")
    ap.add_argument("--init-stddev", type=float, default=1.0)
    ap.add_argument("--min-seconds", dest="min_seconds", type=float, default=0.3,
                    help="repeat the K-step measurement until this much wall time has been spent (stable mean, clock samples)")
    ap.add_argument("--precision", default=None, choices=["tf32", "bf16"],
                    help="tensor-core operand type: tf32 (fp32 activations/weights, the reference's precision class; "
                         "default) or bf16 (bf16 operands, fp32 master weights)")
    ap.add_argument("--graphs", action="store_true", help="--impl ddp: capture the whole DDP step in a CUDA graph")
    ap.add_argument("--compile", action="store_true", help="--impl ddp: wrap the model in torch.compile")
    return ap.parse_args()


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def main():
    a = parse()
    rank, world, local = dist_env()
    if a.impl == "reference":
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": "reference = 6 TensorFlow-1.x scripts (tf.app.flags, "
                              "tf.train.Supervisor, tensorflow.examples.tutorials.mnist); tensorflow is not installed, "
                              "not in /opt/wheelhouse and there is no network; /root/reference has no setup.py/pyproject "
                              "so pip cannot install it either"}))
        return 0
    import torch
    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device", "impl": a.impl}))
        return 1
    if a.gpus != world:
        if world == 1 and a.gpus > 1:
            # convenience: re-launch ourselves under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(a.gpus),
                   "--master-addr", "127.0.0.1", "--master-port", str(29400 + os.getpid() % 500), __file__] + sys.argv[1:]
            return subprocess.call(cmd)
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (a.gpus, world))
    torch.cuda.set_device(local)
    if a.impl == "ddp":
        from baseline.ddp_baseline import run_ddp
        res = run_ddp(a, rank, world, local)
    else:
        from tensorflow_distributed_b200.benchmark import run_ours
        res = run_ours(a, rank, world, local)
    if rank == 0 and res is not None:
        print(json.dumps(res))
    return 0


if __name__ == "__main__":
    sys.exit(main())
