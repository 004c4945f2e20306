"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv`) per kernel: launches, total time, share.

    python tools/summarize_launches.py gpurun_out/launches.csv > profiles/rNN_launches_summary.txt
"""
import csv
import collections
import re
import sys


def short(name):
    name = re.sub(r"\(.*", "", name)  # drop the argument list
    name = name.replace("sfb::", "").replace("void ", "")
    return name[:90]


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        ns = float(r["Metric Value"].replace(",", ""))
        if r.get("Metric Unit") == "us":
            ns *= 1e3
        elif r.get("Metric Unit") == "ms":
            ns *= 1e6
        rows.append((short(r["Kernel Name"]), ns))
    tot = sum(ns for _, ns in rows)
    agg = collections.defaultdict(lambda: [0, 0.0])
    for k, ns in rows:
        agg[k][0] += 1
        agg[k][1] += ns
    print(f"# {path}: {len(rows)} launches, {tot / 1e6:.3f} ms summed kernel time (ncu: serialised, cold caches)")
    print(f"{'kernel':90s} {'n':>5s} {'ms':>9s} {'share':>7s}")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:90s} {n:5d} {ns / 1e6:9.3f} {100 * ns / tot:6.2f}%")


if __name__ == "__main__":
    main(sys.argv[1])
