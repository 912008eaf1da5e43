"""Compact per-kernel SASS listings for profiles/sass/ (address + instruction, encodings dropped):
    python scripts/sass_dump.py            # rewrites the files named in KERNELS from the in-tree library"""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tensorflow_distributed_b200", "_native", "libtfd_b200.so")
KERNELS = {  # output file -> regex on the mangled name (templates: ...IL<type>...; TF = 1 is the tf32 arm)
    # the CNN step, tf32 arm (bench default) and bf16 arm
    "conv1_fwd_f32.sass": r"conv1_fwd_kernelIfE",
    "conv1_fwd_bf16.sass": r"conv1_fwd_kernelI13__nv_bfloat16E",
    "conv2_fwd_tf32.sass": r"conv2_tc_kernelILb1ELi1E",
    "conv2_fwd_bf16.sass": r"conv2_tc_kernelILb1ELi0E",
    "conv2_dgrad_tf32.sass": r"conv2_tc_kernelILb0ELi1E",
    "conv2_dgrad_bf16.sass": r"conv2_tc_kernelILb0ELi0E",
    "conv2_box_fwd_tf32.sass": r"conv2_box_kernelILb1ELi1E",
    "conv2_box_dgrad_tf32.sass": r"conv2_box_kernelILb0ELi1E",
    "conv2_box_dgrad_bf16.sass": r"conv2_box_kernelILb0ELi0E",
    "conv2_wgrad_tf32.sass": r"conv2_wgrad_tc_kernelILi1E",
    "conv2_wgrad_bf16.sass": r"conv2_wgrad_tc_kernelILi0E",
    "gemm_fc1_fwd_kA_mnB_bn64_tf32.sass": r"gemm_tc_kernelILi0ELi1ELi64ELi1E",
    "gemm_fc1_fwd_kA_mnB_bn64_bf16.sass": r"gemm_tc_kernelILi0ELi1ELi64ELi0E",
    "gemm_fc1_dgrad_kA_kB_bn64_tf32.sass": r"gemm_tc_kernelILi0ELi0ELi64ELi1E",
    "gemm_fc1_dgrad_kA_kB_bn64_bf16.sass": r"gemm_tc_kernelILi0ELi0ELi64ELi0E",
    "gemm_fc1_wgrad_mnA_mnB_bn128_tf32.sass": r"gemm_tc_kernelILi1ELi1ELi128ELi1E",
    "gemm_fc1_wgrad_mnA_mnB_bn128_bf16.sass": r"gemm_tc_kernelILi1ELi1ELi128ELi0E",
    "head_rows_f32.sass": r"head_rows_kernelIfE",
    "head_cols_f32.sass": r"head_cols_kernelIfE",
    "pool2_bwd_f32.sass": r"pool2_bwd_kernelIfE",
    "conv1_wgrad_f32.sass": r"conv1_wgrad_kernelIfE",
    # parameter-server kernels
    "adam_apply.sass": r"adam_apply_kernel",
    "signal.sass": r"signal_kernel",
    "gate.sass": r"gate_kernel",
    "decide.sass": r"decide_kernel",
    "pull.sass": r"pull_kernel",
    "push.sass": r"push_kernel",
    "nvls_apply.sass": r"nvls_apply_kernel",
    "nvls_signal.sass": r"nvls_signal_kernel",
    "sum_planes_noalloc.sass": r"sum_planes_kernel",
    "random_normal.sass": r"random_normal_kernel",
}

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
blocks, cur = {}, None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        blocks[cur] = []
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if cur and m:
        blocks[cur].append("%s  %s" % (m.group(1), m.group(2).strip()))
os.makedirs(os.path.join(ROOT, "profiles", "sass"), exist_ok=True)
for old in os.listdir(os.path.join(ROOT, "profiles", "sass")):
    if old.endswith(".sass") and old not in KERNELS:
        os.remove(os.path.join(ROOT, "profiles", "sass", old))
for fname, pat in KERNELS.items():
    fn = next(f for f in blocks if re.search(pat, f))
    with open(os.path.join(ROOT, "profiles", "sass", fname), "w") as f:
        f.write("\nFatbin elf code:\n================\narch = sm_100a\nFunction : %s\n" % fn)
        f.write("\n".join(blocks[fn]) + "\n")
    print(fname, len(blocks[fn]), "instructions")
