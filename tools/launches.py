"""Print kernel name + device time from an ncu launch-list csv (gpu__time_duration.sum)."""
import csv
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")))
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
n = int(sys.argv[2]) if len(sys.argv) > 2 else 12
for r in rows[1:1 + n]:
    print("%-80s %8.1f us" % (r[ki][:80], float(r[vi]) / 1e3))
