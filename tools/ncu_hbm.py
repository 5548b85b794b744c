#!/usr/bin/env python
"""Summarise an `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,... --csv` capture of
`python tools/prof_step.py --ngrid 256` (bootstrap step + one measured step = 2 passes of every stage): DRAM bytes per
kernel and, per stage, measured DRAM bytes against the algorithmic bytes of DESIGN.md section 4.
usage: ncu_hbm.py hbm_kernels.csv [numnodes] > profiles/rN_hbm_kernels_*.txt"""
import csv, sys, collections
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")) if r]
numnodes = float(sys.argv[2]) if len(sys.argv) > 2 else 7075020.0
h = rows[0]
ki, mi, vi, ui, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit"), h.index("ID")
per = {}
for r in rows[1:]:
    if len(r) <= vi: continue
    k = (r[ii], r[ki].split("(")[0][:48])
    v = float(r[vi].replace(",", "")); u = r[ui]
    if r[mi].startswith("dram__bytes"): v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    if r[mi] == "gpu__time_duration.sum": v *= {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3}.get(u, 1e-6)
    per.setdefault(k, {})[r[mi]] = v
agg = collections.OrderedDict()
for (i, k), m in per.items():
    a = agg.setdefault(k, [0, 0.0, 0.0, 0.0, 0.0])
    a[0] += 1; a[1] += m.get("gpu__time_duration.sum", 0); a[2] += m.get("dram__bytes_read.sum", 0); a[3] += m.get("dram__bytes_write.sum", 0)
    a[4] += m.get("lts__t_sector_hit_rate.pct", 0)
print("# ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sector_hit_rate.pct --clock-control none: python tools/prof_step.py --ngrid 256")
print("# (256^3, Nmesh 512: Barnes-Hut bootstrap step + one measured step = 2 passes of every stage; sums over both; times are under the profiler (cold, serialised): use the BYTES, and bench.py's stage timers for the times)")
print("%-50s %5s %9s %10s %10s %9s %7s" % ("kernel", "n", "ms", "read MB", "write MB", "GB/s", "L2hit%"))
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-50s %5d %9.3f %10.1f %10.1f %9.0f %7.1f" % (k, a[0], a[1], a[2] / 1e6, a[3] / 1e6, (a[2] + a[3]) / 1e9 / (a[1] * 1e-3) if a[1] > 0 else 0, a[4] / a[0]))
n = 16777216; nm3 = 512 ** 3
print("\n# algorithmic bytes per pass (SURVEY 8d / DESIGN 4) vs measured DRAM bytes per pass (= sums above / 2):")
def line(name, alg, keys):
    tot = sum(agg[k][2] + agg[k][3] for k in agg if any(s in k for s in keys)) / 2
    ms = sum(agg[k][1] for k in agg if any(s in k for s in keys)) / 2
    print("#   %-28s algorithmic %7.1f MB   DRAM %7.1f MB   ratio %.2f   (%.3f ms under ncu)" % (name, alg / 1e6, tot / 1e6, tot / alg, ms))
line("PH keys", 32.0 * n, ["sq_keys_kernel"])
line("radix sort (8 passes)", 8 * 32.0 * n, ["sq_sort_hist", "sq_sort_scatter", "sq_scan"])
line("gather", 80.0 * n, ["sq_gather_kernel"])
line("tree build (split+moments)", 36.0 * n + 80.0 * numnodes + 64.0 * numnodes, ["sq_tree_split", "sq_tree_moments"])
line("CIC paint", 28.0 * n + 16.0 * nm3, ["sq_pm_paint"])
line("transfer", 32.0 * nm3 / 2, ["sq_pm_transfer"])
line("gradient", 32.0 * nm3, ["sq_pm_gradient"])
line("CIC readout", 56.0 * n + 32.0 * nm3, ["sq_pm_readout"])
