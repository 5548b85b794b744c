import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
import shenqi_b200 as sq
from helpers import *
from test_gpu_parity import make_params, oracle_walk_for, to_oracle_par
ctx = sq.Context(0)
box = 8.0; n = 16 ** 3
pos = random_clustered(n, box, 4); mass = np.ones(n, np.float32)
for g in (32, 16, 8):
    ctx.set_group_width(g)
    P = make_params(box, 48, n, TreeUseBH=1, TreeRcut=7.0)
    ctx.set_params(P)
    ctx.upload_particles(pos, mass); ctx.ph_sort(); ctx.tree_build()
    a, _, st = ctx.grav_short_tree()
    o, _, _, ot = oracle_walk_for(pos, mass, box, P, np.zeros(n))
    pa, _ = o_pair(pos, mass, to_oracle_par(P))
    e_ours, e_ref = rel_err(a, pa), rel_err(o, pa)
    d = rel_err(a, o)
    print(g, "int", st.interactions, ot, "mean", e_ours.mean(), e_ref.mean(), "max", e_ours.max(), e_ref.max(), "ours-vs-oracle median %.3e p99.9 %.3e max %.3e argmax %d" % (np.median(d), np.percentile(d, 99.9), d.max(), d.argmax()))
    i = d.argmax(); print(a[i], o[i], pa[i])
