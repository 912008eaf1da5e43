"""profiles/ncu_step_kernels.md from the ncu reports of one training step per precision arm:

    ncu -i gpurun_out/prof_step_tf32.ncu-rep --page raw --csv > /tmp/tf32.csv   (same for bf16)
    python scripts/ncu_step_table.py /tmp/tf32.csv /tmp/bf16.csv > profiles/ncu_step_kernels.md

One row per kernel of the step: ncu duration (cold caches, kernels serialised -- compare shares, not absolutes), tensor
pipe, occupancy, DRAM traffic, and the fraction of the kernel's roofline: ideal = max(algorithmic bytes / measured HBM
copy bandwidth, FLOPs / measured GEMM peak) with the peaks of MEASURED_PEAKS.json (tf32 tensor peak = half of bf16's)."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    PK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    HBM, TF_BF16, SRC = PK["hbm_gbs"] * 1e9, PK["bf16_tflops"] * 1e12, "measured"
except Exception:  # noqa: BLE001
    HBM, TF_BF16, SRC = 6.65e12, 1.59e15, "fallback"
B = 128
# algorithmic work per kernel at B = 128: (bytes fp32-arm, bytes bf16-arm, tensor FLOPs, CUDA-core FLOPs)
WORK = {
    "conv1_fwd_kernel": (B * 784 * 4 + B * 196 * 32 * 5, B * 784 * 4 + B * 196 * 32 * 3, 0, 2 * B * 784 * 32 * 25),
    "conv2_tc_kernel<1": (B * 196 * 32 * 4 + 204800 + B * 49 * 64 * 5, B * 196 * 32 * 2 + 102400 + B * 49 * 64 * 3, 2 * B * 196 * 64 * 800, 0),
    "conv2_tc_kernel<0": (B * 196 * 64 * 4 + 204800 + 2 * B * 196 * 32 * 4, B * 196 * 64 * 2 + 102400 + B * 196 * 32 * 4, 2 * B * 196 * 64 * 800, 0),
    "conv2_box_kernel<1": (B * 196 * 32 * 4 + 204800 + B * 49 * 64 * 5, B * 196 * 32 * 2 + 102400 + B * 49 * 64 * 3, 2 * B * 196 * 64 * 800, 0),
    "conv2_box_kernel<0": (B * 196 * 64 * 4 + 204800 + 2 * B * 196 * 32 * 4, B * 196 * 64 * 2 + 102400 + B * 196 * 32 * 4, 2 * B * 196 * 64 * 800, 0),
    "conv2_wgrad_tc_kernel": (B * 196 * 96 * 4 + 204800, B * 196 * 96 * 2 + 204800, 2 * B * 196 * 64 * 800, 0),
    "gemm_tc_kernel<0, 1": (3136 * 1024 * 4 + B * 3136 * 4 + 8 * B * 1024 * 4, 3136 * 1024 * 2 + B * 3136 * 2 + 8 * B * 1024 * 4, 2 * B * 1024 * 3136, 0),
    "gemm_tc_kernel<0, 0": (3136 * 1024 * 4 + B * 1024 * 4 + 3 * B * 3136 * 4, 3136 * 1024 * 2 + B * 1024 * 2 + 3 * B * 3136 * 4, 2 * B * 1024 * 3136, 0),
    "gemm_tc_kernel<1, 1": (3136 * 1024 * 4 + B * 4160 * 4, 3136 * 1024 * 4 + B * 4160 * 2, 2 * B * 1024 * 3136, 0),
    "head_rows_kernel": (8 * B * 1024 * 4 + 2 * B * 1024 * 4, 8 * B * 1024 * 4 + 2 * B * 1024 * 2, 0, 4 * B * 1024 * 10),
    "head_cols_kernel": (2 * B * 1024 * 4 + 45000, 2 * B * 1024 * 2 + 45000, 0, 2 * B * 1024 * 11),
    "pool2_bwd_kernel": (3 * B * 3136 * 4 + B * 3136 * 5 + B * 196 * 64 * 4, 3 * B * 3136 * 4 + B * 3136 * 3 + B * 196 * 64 * 2, 0, 0),
    "conv1_wgrad_kernel": (B * 784 * 4 + B * 196 * 32 * 13, B * 784 * 4 + B * 196 * 32 * 7, 0, 2 * B * 196 * 32 * 25),
    "adam_apply_kernel": (3274634 * 4 * 7, 3274634 * 4 * 7 + 3262464 * 2, 0, 0),
}


def table(path, arm):
    rows = list(csv.reader(open(path)))
    hdr, data = rows[0], rows[2:]
    col = {n: hdr.index(n) for n in ("Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__registers_per_thread",
                                     "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
                                     "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum",
                                     "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed") if n in hdr}
    units = rows[1]
    peak = TF_BF16 * (0.5 if arm == "tf32" else 1.0)
    print("\n## %s arm\n" % arm)
    print("| kernel | ncu us | grid | regs | tensor pipe %% | warps active %% | DRAM rd+wr MB | SM %% | ideal us | fraction of roofline (of %s) |" % SRC)
    print("|---|---|---|---|---|---|---|---|---|---|")
    seen = set()
    for r in data:
        name = r[col["Kernel Name"]].replace("<unnamed>::", "").replace("void ", "").split("(")[0]
        if name in seen:
            continue
        seen.add(name)
        us = float(r[col["gpu__time_duration.sum"]].replace(",", ""))

        def mb(c):
            v, u = float(r[col[c]].replace(",", "")), units[col[c]]
            return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1.0)
        key = next((k for k in WORK if name.startswith(k)), None)
        ideal = frac = ""
        if key:
            b32, b16, tfl, cfl = WORK[key]
            t = max((b32 if arm == "tf32" else b16) / HBM, tfl / peak, cfl / 72e12) * 1e6
            ideal, frac = "%.1f" % t, "%.2f" % (t / us)
        print("| `%s` | %.1f | %s | %s | %.1f | %.1f | %.1f | %.1f | %s | %s |" % (
            name[:44], us, r[col["launch__grid_size"]].split(".")[0], r[col["launch__registers_per_thread"]].split(".")[0],
            float(r[col["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]]),
            float(r[col["sm__warps_active.avg.pct_of_peak_sustained_active"]]),
            mb("dram__bytes_read.sum") + mb("dram__bytes_write.sum"),
            float(r[col["sm__throughput.avg.pct_of_peak_sustained_elapsed"]]), ideal, frac))


print("# ncu --set full of one CNN training step (B = 128, 1 GPU, round 2 final kernels)\n")
print("Capture dates: tf32 arm after the one-box conv2 kernels, the Adam fence changes and the dgrad-first capture order, but BEFORE the\n"
      "conv1-wgrad rewrite (conflict-free image rows + two-level final sum: 24.2 -> 16.7 us in the step graph, `profiles/step_timeline.md`);\n"
      "bf16 arm one change set earlier (five-box conv2 forward, one-box dgrad not yet default).  The GPU budget of the round ended before a re-capture.\n")
print("Captured with `ncu --set full --clock-control none --import-source on -s 60 -c 12 python bench.py --steps 2 ...` after the same\n"
      "command had exited 0 without ncu.  ncu serialises the kernels and starts each one on cold caches, so durations are\n"
      "30-60 %% above the in-graph durations of `profiles/step_timeline.md`; the columns to read are the shares and the\n"
      "pipe / occupancy numbers.  Roofline inputs: HBM %.0f GB/s, bf16 %.0f TFLOP/s (tf32: half), CUDA-core fp32 72 TFLOP/s;\n"
      "ideal = max(algorithmic bytes / HBM, FLOPs / peak).  Every kernel but Adam is far from its roofline because it does\n"
      "1-2 us of ideal work inside a 6-20 us kernel: the step is latency / L2-traffic bound (DESIGN.md, last section)." % (HBM / 1e9, TF_BF16 / 1e12))
for path in sys.argv[1:]:
    table(path, "tf32" if "tf32" in path else "bf16")
