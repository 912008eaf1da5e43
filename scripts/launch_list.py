"""Print the kernels of the last full training step from an `ncu --metrics gpu__time_duration.sum --csv` log."""
import csv
import sys

lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
seq = [(r["Kernel Name"], float(r["Metric Value"])) for r in csv.DictReader(lines)]
first = sys.argv[2] if len(sys.argv) > 2 else "zero_multi"
idx = [i for i, (n, _) in enumerate(seq) if first in n]
a, b = idx[-3], idx[-2]
tot = 0.0
for n, t in seq[a:b]:
    print("%-64s %8.1f us" % (n.replace("<unnamed>::", "")[:64], t / 1000))
    tot += t
print("%-64s %8.1f us  (%d kernels)" % ("total", tot / 1000, b - a))
