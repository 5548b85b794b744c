#!/usr/bin/env python
"""Regenerates tests/golden/walk_ref.npz from the reference's OWN code (oracle/_ref/libshenqi_ref.so, built from
/root/reference by oracle/Makefile; run in the build container only).  Contents: a seeded 2048-particle clustered
set, the reference's PH keys, sorted ID order, tree cell list (canonical form), grav_short_tree accelerations for
the Barnes-Hut call and the relative-criterion call (libgadget/tests/test_gravity.cpp:237-240), total interaction
counts of both, and the grav_short_pair direct sum."""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H

box, n, nmesh = 8.0, 2048, 48
pos = H.random_clustered(n, box, 12345)
mass = (1 + np.random.default_rng(99).random(n)).astype(np.float32)
ids = np.arange(1, n + 1, dtype=np.uint64)
meansep = box / round(n ** (1 / 3.))
par = H.default_params(box, nmesh, n, TreeUseBH=1, TreeRcut=7.0)
R = H.ref()
assert R is not None, "build oracle/_ref first (make -C oracle)"
keys = np.zeros(n, np.uint64)
R.ref_peano_keys(H.ptr(pos), H.c_i64(n), H.c_dbl(box), H.ptr(keys))
H.r_load(pos, mass, box, ids)
R.ref_sort_particles()
spos, smass, sid = H.r_sorted_particles()
nodes, first, _ = H.r_tree(0.9)
cells, parts, moms = H.canonical_tree(nodes, first, sid.astype(np.int64), box)
_, t0 = H.r_counts(par, meansep)
a0, _ = H.r_walk(par, meansep)
par2 = dict(par, TreeUseBH=0)
R.ref_set_accels(H.ptr(a0), H.ptr(np.zeros((n, 3))))
_, t1 = H.r_counts(par2, meansep)
a1, _ = H.r_walk(par2, meansep, fulltree=a0, gravpm=np.zeros((n, 3)))
pair, _ = H.r_pair(par, meansep)
inv = (sid - 1).astype(np.int64)          # sorted position -> input index
out = dict(pos=pos, mass=mass, keys=keys, sorted_ids=sid, cells=cells, leaf_parts=parts, moments=moms,
           acc_bh=np.zeros((n, 3)), acc_rel=np.zeros((n, 3)), acc_pair=np.zeros((n, 3)), count_bh=np.int64(t0), count_rel=np.int64(t1),
           softening=np.float64(par["GravitySoftening"]))
out["acc_bh"][inv] = a0; out["acc_rel"][inv] = a1; out["acc_pair"][inv] = pair
np.savez_compressed(os.path.join(HERE, "walk_ref.npz"), **out)
print("wrote walk_ref.npz: %d particles, %d cells, counts %d / %d" % (n, len(cells), t0, t1))
