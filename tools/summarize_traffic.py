"""Per-kernel DRAM traffic of one step from an ncu launch list taken with
--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum  ->  text table + JSON (bench.py reads the
JSON for `roofline.traffic`).
    python tools/summarize_traffic.py gpurun_out/traffic.csv profiles/rNN_traffic.json > profiles/rNN_traffic.txt
"""
import collections
import csv
import json
import re
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def short(name):
    return re.sub(r"\(.*", "", name).replace("sfb::", "").replace("void ", "")[:70]


def main(path, out_json):
    lines = [l for l in open(path, newline="") if l.startswith('"')]
    per = collections.defaultdict(lambda: collections.defaultdict(float))
    for r in csv.DictReader(lines):
        v = float(r["Metric Value"].replace(",", "")) * UNIT.get(r["Metric Unit"], 1.0)
        per[(r["ID"], short(r["Kernel Name"]))][r["Metric Name"]] += v
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for (_, k), m in per.items():
        a = agg[k]
        a[0] += 1
        a[1] += m.get("gpu__time_duration.sum", 0.0)
        a[2] += m.get("dram__bytes_read.sum", 0.0)
        a[3] += m.get("dram__bytes_write.sum", 0.0)
    print(f"# {path}: DRAM bytes per kernel over one step (ncu, --clock-control none, caches flushed per launch)")
    print(f"{'kernel':70s} {'n':>4s} {'ms':>8s} {'read GB':>9s} {'write GB':>9s} {'GB/s':>7s}")
    out = {}
    for k, (n, ms, rd, wr) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:70s} {n:4d} {ms:8.3f} {rd / 1e9:9.3f} {wr / 1e9:9.3f} {(rd + wr) / ms / 1e6 if ms else 0:7.0f}")
        out[k] = {"launches": n, "ms": ms, "dram_read_bytes": rd, "dram_write_bytes": wr}
    tot = [sum(v[i] for v in agg.values()) for i in range(4)]
    print(f"{'TOTAL':70s} {tot[0]:4d} {tot[1]:8.3f} {tot[2] / 1e9:9.3f} {tot[3] / 1e9:9.3f}")
    json.dump(out, open(out_json, "w"), indent=1)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
