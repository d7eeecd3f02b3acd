#!/usr/bin/env python3
"""Per-kernel table from `ncu -i X.ncu-rep --page raw --csv`: time, DRAM bytes, hit rates, occupancy."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h, units = rows[0], rows[1]
cols = [("gpu__time_duration.sum", "ms"), ("dram__bytes_read.sum", "rdGB"), ("dram__bytes_write.sum", "wrGB"),
        ("lts__t_sector_hit_rate.pct", "L2hit%"), ("l1tex__t_sector_hit_rate.pct", "L1hit%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid")]
kn = h.index("Kernel Name")
def conv(v, u):
    v = float(v.replace(",", ""))
    if u in ("us", "usecond"): v /= 1e3
    if u in ("ns", "nsecond"): v /= 1e6
    if u == "Mbyte": v /= 1e3
    if u == "Kbyte": v /= 1e6
    if u == "byte": v /= 1e9
    return v
print("%-28s" % "kernel" + "".join("%9s" % c[1] for c in cols) + "   DRAM GB/s")
for r in rows[2:]:
    vals = [conv(r[h.index(c)], units[h.index(c)]) for c, _ in cols]
    bw = (vals[1] + vals[2]) / (vals[0] / 1e3)
    print("%-28s" % r[kn].replace("void ", "")[:27] + "".join("%9.3f" % v if v < 1000 else "%9d" % v for v in vals) + "   %8.0f" % bw)
