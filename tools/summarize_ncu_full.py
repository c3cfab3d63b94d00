"""Dev tool: per-launch summary of an `ncu -i X.ncu-rep --page raw --csv` dump (the text files under profiles/).
usage: python tools/summarize_ncu_full.py raw.csv "header line" > profiles/<tag>_<kernel>_ncu_full.txt
Also prints, on stderr, the mean dram bytes per launch (the `traffic` figure of bench.py, profiles/ncu_traffic.json)."""
import csv
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, body = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
print("# " + (sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]))
traffic = []
for r in body:
    print("---")
    for k in ("Kernel Name", "Grid Size", "Block Size"):
        print("%-80s %s " % (k, r[col[k]][:110]))
    t = 0.0
    for k in KEEP:
        if k in col:
            print("%-80s %s %s" % (k, r[col[k]], units[col[k]]))
            if k.startswith("dram__bytes"):
                t += float(r[col[k]].replace(",", "")) * SCALE.get(units[col[k]], 1.0)
    traffic.append(t)
if traffic:
    sys.stderr.write("launches %d  mean dram bytes per launch %.1f\n" % (len(traffic), sum(traffic) / len(traffic)))
