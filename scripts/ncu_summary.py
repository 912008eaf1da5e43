"""Markdown summary of an ncu report (`ncu -i X.ncu-rep --page raw --csv > X.csv`): one row per kernel launch."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[0], rows[2:]
cols = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "us"), ("sm__cycles_elapsed.max", "SM cycles"),
        ("launch__grid_size", "grid"), ("launch__registers_per_thread", "regs"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
        ("dram__bytes_read.sum", "DRAM rd"), ("dram__bytes_write.sum", "DRAM wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %")]
idx = [(hdr.index(c), n) for c, n in cols if c in hdr]
units = rows[1]
print("| " + " | ".join("%s%s" % (n, (" (" + units[i] + ")") if units[i] and n not in ("kernel",) else "") for i, n in idx) + " |")
print("|" + "---|" * len(idx))
for r in data:
    cells = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = "`" + v.replace("<unnamed>::", "").split("(")[0][:48] + "`"
        else:
            try:
                v = "%.1f" % float(v.replace(",", ""))
            except ValueError:
                pass
        cells.append(v)
    print("| " + " | ".join(cells) + " |")
