#!/usr/bin/env python3
"""Per-kernel device time of one pass of the path from an `ncu --metrics gpu__time_duration.sum --csv`
launch list (launches after the index build's last kernel, k_sak / k_ktab_mark)."""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
h = rows[hi]; kn = h.index('Kernel Name'); mv = h.index('Metric Value')
seq = [(r[kn][:52], float(r[mv].replace(',', '')) / 1e6) for r in rows[hi + 1:] if len(r) > mv]
names = [s[0] for s in seq]
marks = [i for i, nm in enumerate(names) if nm.startswith('k_sak') or nm.startswith('k_ktab_mark')]
i0 = (max(marks) + 1) if marks else 0
while i0 < len(seq) and not seq[i0][0].startswith('k_'): i0 += 1
agg = collections.OrderedDict(); tot = 0
for nm, ms in seq[i0:]:
    agg.setdefault(nm, [0, 0]); agg[nm][0] += ms; agg[nm][1] += 1; tot += ms
for nm, (ms, c) in agg.items():
    print("%8.3f ms %5.1f%% %3d  %s" % (ms, 100 * ms / tot, c, nm))
print("%8.3f ms total" % tot)
