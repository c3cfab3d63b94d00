"""Dev tool: top stalled SASS instructions from `ncu -i X.ncu-rep --page source --csv` (first kernel in the file).
usage: ncu -i rep --page source --csv | python tools/ncu_top_stalls.py [N]"""
import csv
import sys

n = int(sys.argv[1]) if len(sys.argv) > 1 else 25
rows = list(csv.reader(sys.stdin))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
body = []
for r in rows[hi + 1:]:
    if r and r[0] == "Kernel Name":
        break
    if len(r) == len(hdr):
        body.append(r)
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[col["# Samples"]] or 0) for r in body)
print("total samples", tot, "instructions", len(body))
agg = {s: sum(int(r[col[s]] or 0) for r in body) for s in stalls}
print("by reason:", ", ".join("%s=%d" % (k[6:], v) for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
for r in sorted(body, key=lambda r: -int(r[col["# Samples"]] or 0))[:n]:
    why = sorted(((int(r[col[s]] or 0), s[6:]) for s in stalls), reverse=True)[:3]
    print("%6d  %-70s %s" % (int(r[col["# Samples"]]), r[col["Source"]].strip()[:70], " ".join("%s=%d" % (b, a) for a, b in why if a)))
