"""List the SASS instructions holding most pc-sampling samples of one launch.
    ncu -i X.ncu-rep --page source --csv --launch-skip K --launch-count 1 --print-source sass > src.csv
    python tools/hot_sass.py src.csv [min_pct]
"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.7
hdr = next(r for r in rows if "Source" in r and "# Samples" in r)
data = [r for r in rows if len(r) == len(hdr) and r is not hdr and r[hdr.index("# Samples")].isdigit()]
ia, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
tot = sum(int(r[isamp]) for r in data)
print("total samples", tot, "instructions", len(data))
for i, r in enumerate(data):
    s = int(r[isamp])
    if s > tot * thr / 100:
        print(f"{i:5d} {s:7d} {100 * s / tot:5.1f}%  exec {r[iex]:>9s}  {r[ia][:110]}")
