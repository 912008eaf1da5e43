"""Compact per-kernel SASS listings for profiles/sass/ (address + instruction, encodings dropped):
    python scripts/sass_dump.py            # rewrites the files named in KERNELS from the in-tree library"""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tensorflow_distributed_b200", "_native", "libtfd_b200.so")
KERNELS = {  # output file -> regex on the mangled name
    "adam_apply.sass": r"adam_apply_kernel",
    "conv1_fwd.sass": r"conv1_fwd_kernel",
    "conv2_dgrad.sass": r"conv2_tc_kernelILb0",
    "conv2_fwd.sass": r"conv2_tc_kernelILb1",
    "conv2_wgrad.sass": r"conv2_wgrad_tc_kernel",
    "gemm_tc_fwd_kmajorA_mnmajorB_bn64.sass": r"gemm_tc_kernelILi0ELi1ELi64",
    "gemm_tc_wgrad_mnmajor_bn128.sass": r"gemm_tc_kernelILi1ELi1ELi128",
    "pull.sass": r"pull_kernel",
    "push.sass": r"push_kernel",
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
for fname, pat in KERNELS.items():
    fn = next(f for f in blocks if re.search(pat, f))
    with open(os.path.join(ROOT, "profiles", "sass", fname), "w") as f:
        f.write("\nFatbin elf code:\n================\narch = sm_100a\nFunction : %s\n" % fn)
        f.write("\n".join(blocks[fn]) + "\n")
    print(fname, len(blocks[fn]), "instructions")
