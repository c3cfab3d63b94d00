"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name, launches, total and share.
usage: python tools/summarize_launches.py gpurun_out/<tag>_launches.csv [skip_first_n] > profiles/<tag>_launches_summary.txt"""
import collections
import csv
import re
import sys

path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = []
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r["Metric Name"] != "gpu__time_duration.sum":
        continue
    rows.append((r["Kernel Name"], float(r["Metric Value"]) / 1e3, r["Grid Size"], r["Block Size"]))
rows = rows[skip:]
agg = collections.OrderedDict()
for name, us, grid, block in rows:
    short = re.sub(r"\(.*", "", name).replace("void ", "").replace("(anonymous namespace)::", "")
    short = re.sub(r"^at::native::.*?(\w+_kernel).*", r"torch:\1", short)
    d = agg.setdefault(short, [0, 0.0])
    d[0] += 1
    d[1] += us
tot = sum(v[1] for v in agg.values())
print("# %s: %d launches, %.1f us summed device time (cold-cache, serialised under ncu: compare SHARES)" % (path, len(rows), tot))
print("%-64s %7s %11s %7s %9s" % ("kernel", "count", "total_us", "share", "avg_us"))
for k, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-64s %7d %11.1f %6.1f%% %9.2f" % (k[:64], c, us, 100 * us / tot, us / c))
