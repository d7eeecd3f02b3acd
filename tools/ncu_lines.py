#!/usr/bin/env python3
"""Aggregate an `ncu --page source --print-source cuda,sass --csv` dump per CUDA source line."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
fname = None; hdr = None; line = None; text = {}
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
for r in rows:
    if len(r) == 2 and r[0] == "File Name":
        fname = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No":
        hdr = r; iS = hdr.index("# Samples"); iI = hdr.index("Instructions Executed")
        iL = hdr.index("L2 Theoretical Sectors Global"); iT = hdr.index("Thread Instructions Executed"); continue
    if hdr is None or len(r) < len(hdr): continue
    if r[0] != "":
        line = (fname, int(r[0])); text[line] = r[1]
        continue
    if r[2] in ("...", ""): continue
    try:
        a = agg[line]; a[0] += int(r[iS] or 0); a[1] += int(r[iI] or 0); a[2] += int(r[iL] or 0); a[3] += int(r[iT] or 0)
    except ValueError:
        pass
ts = sum(a[0] for a in agg.values()); ti = sum(a[1] for a in agg.values()); tl = sum(a[2] for a in agg.values())
print("samples %d  warp-inst %d  L2 sectors %d" % (ts, ti, tl))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% smp %5.1f%% inst %5.1f%% l2  thr/inst %4.1f  %s:%d  %s" % (
        100 * a[0] / max(ts, 1), 100 * a[1] / max(ti, 1), 100 * a[2] / max(tl, 1), a[3] / max(a[1], 1), k[0], k[1], text[k].strip()[:90]))
