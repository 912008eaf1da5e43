"""Per-kernel SASS evidence: which Blackwell-native instructions each kernel of libtfd_b200.so contains."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "tensorflow_distributed_b200/_native/libtfd_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pats = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMAPF", "LDTM", "SYNCS", "UTCATOM", "LDGMC", "RED.E", "REDG", "ATOMG", "LDG", "STG", "HMMA", "SHFL", "MEMBAR", "ERRBAR"]
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(anonymous namespace\)::", "", cur).split("(")[0]
        counts[cur] = collections.Counter()
        continue
    if cur:
        for p in pats:
            if re.search(r"\b" + re.escape(p), line):
                counts[cur][p] += 1
print("| kernel | " + " | ".join(pats) + " |")
print("|---|" + "---|" * len(pats))
for k, c in counts.items():
    print("| `%s` | " % k[:70] + " | ".join(str(c[p]) if c[p] else "" for p in pats) + " |")
