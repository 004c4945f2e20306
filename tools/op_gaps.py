"""Per-shape gap between the measured conv launches of one step (SFB_DUMP_OPS file) and their roofline-ideal time.
    python tools/op_gaps.py gpurun_out/ops.txt [batch] [top]
"""
import collections
import re
import sys

N = int(sys.argv[2]) if len(sys.argv) > 2 else 64
TOP = int(sys.argv[3]) if len(sys.argv) > 3 else 40
PEAK_TF, PEAK_GB = 1422.7e12, 6.2e12
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
kinds = collections.defaultdict(lambda: [0.0, 0.0])
for l in open(sys.argv[1]):
    m = re.match(r"\s*([\d.]+) ms\s+([\d.]+) TF/s (\S+)\s+(.*)", l)
    if not m:
        continue
    ms, tf, desc = float(m.group(1)), float(m.group(2)), m.group(4)
    d = re.match(r"(\w+) in\((\d+),(\d+),(\d+)\)x(\d+) out\((\d+),(\d+),(\d+)\)x(\d+) k", desc)
    if not d:
        continue
    v = list(map(int, d.groups()[1:]))
    byt = N * 2 * (v[0] * v[1] * v[2] * v[3] + v[4] * v[5] * v[6] * v[7])
    ideal = max(tf * ms * 1e9 / PEAK_TF, byt / PEAK_GB) * 1e3
    a = agg[desc]
    a[0] += 1; a[1] += ms; a[2] += ideal
    kinds[d.group(1)][0] += ms; kinds[d.group(1)][1] += ideal
for k, (ms, ideal) in kinds.items():
    print(f"{k:6s} {ms:7.3f} ms  ideal {ideal:7.3f}")
for desc, (n, ms, ideal) in sorted(agg.items(), key=lambda kv: -(kv[1][1] - kv[1][2]))[:TOP]:
    print(f"{n:3d} {ms:7.3f} ideal {ideal:7.3f} gap {ms - ideal:7.3f}  {desc}")
