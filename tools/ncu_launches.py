#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: launches, total ms and share per kernel.
usage: ncu_launches.py launches.csv [header line ...] > profiles/rN_launches_*.txt"""
import csv, sys, collections
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")) if r]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= vi: continue
    try: v = float(r[vi].replace(",", ""))
    except ValueError: continue
    ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
    k = r[ki].split("(")[0][:60]
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += ms
tot = sum(a[1] for a in agg.values()); n = sum(a[0] for a in agg.values())
for h in sys.argv[2:]: print("# " + h)
print("# per-launch times are cold-cache and serialised; shares are what matters.  total %.1f ms over %d launches" % (tot, n))
print("%-62s %6s %10s %7s" % ("kernel", "n", "ms", "share"))
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-62s %6d %10.3f %6.2f%%" % (k, a[0], a[1], 100 * a[1] / tot))
