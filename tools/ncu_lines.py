#!/usr/bin/env python
"""Attribute the per-SASS-instruction counters of an `ncu --page source --csv` export to CUDA source lines.
The CSV has no line column, so the n-th SASS row is joined with the n-th instruction of
`nvdisasm -g <cubin>` for the same kernel (both are in address order).
usage: ncu_lines.py src.csv walk.sass <mangled-kernel-name-substring> [top]"""
import csv, re, sys, collections

src_csv, sass, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
rows = list(csv.reader(open(src_csv)))
h = rows[1]
ii, sm, ti = h.index("Instructions Executed"), h.index("# Samples"), h.index("Thread Instructions Executed")
data = []
for r in rows[2:]:
    try:
        data.append((int(r[ii] or 0), int(r[sm] or 0), int(r[ti] or 0), r[1].strip()))
    except Exception:
        pass
# line info per instruction of the kernel
lines, cur, inside = [], None, False
for ln in open(sass):
    if ln.startswith(".text."):
        inside = kern in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        lines.append(cur)
print("csv rows %d, sass instructions %d" % (len(data), len(lines)))
n = min(len(data), len(lines))
agg = collections.defaultdict(lambda: [0, 0, 0])
for (a, b, c, _), l in zip(data[:n], lines[:n]):
    agg[l][0] += a; agg[l][1] += b; agg[l][2] += c
tot = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
srcs = {}
def text(l):
    if l is None:
        return ""
    f, k = l
    if f not in srcs:
        import glob
        g = glob.glob("/root/repo/shenqi_b200/csrc/" + f)
        srcs[f] = open(g[0]).read().split("\n") if g else []
    return srcs[f][k - 1].strip()[:90] if k - 1 < len(srcs[f]) else ""
print("total warp instructions %d" % tot)
for l, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.2f%% inst %5.2f%% smp thr %4.1f  %s:%s  %s" % (100 * v[0] / tot, 100 * v[1] / max(ts, 1), v[2] / max(v[0], 1), l[0] if l else "?", l[1] if l else "?", text(l)))
