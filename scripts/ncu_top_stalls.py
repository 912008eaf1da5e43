"""Summarise an `ncu --page source --csv` dump: the SASS lines with the most stall samples."""
import csv
import sys

path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
hdr = rows[1]
i_src, i_n = hdr.index("Source"), hdr.index("# Samples")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = []
for r in rows[2:]:
    try:
        n = int(r[i_n])
    except Exception:
        continue
    data.append((n, r))
tot = sum(n for n, _ in data)
print("total samples", tot)
for n, r in sorted(data, key=lambda t: -t[0])[:top]:
    st = sorted(((int(r[i] or 0), h) for i, h in stall_cols), reverse=True)[:2]
    print("%6d %5.1f%%  %-70s %s" % (n, 100.0 * n / max(tot, 1), r[i_src][:70], " ".join("%s=%d" % (h[6:], c) for c, h in st if c)))
