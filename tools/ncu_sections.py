#!/usr/bin/env python
"""Like ncu_lines.py, but sums warp instructions over named line ranges of sq_walk.cu.
usage: ncu_sections.py src.csv walk.sass <kernel-substring> name:lo-hi [name:lo-hi ...]"""
import csv, re, sys, collections
src_csv, sass, kern = sys.argv[1:4]
secs = []
for a in sys.argv[4:]:
    n, r = a.split(":"); lo, hi = r.split("-"); secs.append((n, int(lo), int(hi)))
rows = list(csv.reader(open(src_csv)))
h = rows[1]
ii, ti, si = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("# Samples")
data = []
for r in rows[2:]:
    try: data.append((int(r[ii] or 0), int(r[ti] or 0), int(r[si] or 0)))
    except Exception: pass
lines, cur, inside = [], None, False
for ln in open(sass):
    if ln.startswith(".text."):
        inside = kern in ln; continue
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln): lines.append(cur)
agg = collections.OrderedDict((n, [0, 0, 0]) for n, _, _ in secs); agg["other"] = [0, 0, 0]
for (a, t, sm), l in zip(data, lines):
    k = "other"
    if l and l[0] == "sq_walk.cu":
        for n, lo, hi in secs:
            if lo <= l[1] <= hi: k = n; break
    elif l: k = "hdr:" + l[0]; agg.setdefault(k, [0, 0, 0])
    agg[k][0] += a; agg[k][1] += t; agg[k][2] += sm
tot = sum(v[0] for v in agg.values())
print("total warp instructions %d" % tot)
stot = max(1, sum(v[2] for v in agg.values()))
for k, v in agg.items():
    print("%-22s %6.2f%% inst  %6.2f%% samples  lanes/inst %5.1f" % (k, 100.0 * v[0] / tot, 100.0 * v[2] / stot, v[1] / max(1, v[0])))
