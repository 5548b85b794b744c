#!/usr/bin/env python
"""Runs the reference's public grav_short_tree(..., UseGPU=true) inside tests/dropin/_build/libshenqi_dropin.so
(= reference objects + the B200 hook) on a particle set read from an .npz, twice (Barnes-Hut, then relative
criterion) as libgadget/tests/test_gravity.cpp:237-240, and writes the results.  Own process: the library
defines the same globals as oracle/_ref/libshenqi_ref.so."""
import ctypes, os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H

inp, out = sys.argv[1], sys.argv[2]
d = np.load(inp)
pos, mass, box, nmesh = d["pos"], d["mass"], float(d["box"]), int(d["nmesh"])
n = len(pos)
H.REF_SO = os.path.join(HERE, "_build", "libshenqi_dropin.so")
R = H.ref()
assert R is not None, "tests/dropin/_build/libshenqi_dropin.so missing (make -C tests/dropin)"
R.ref_set_use_gpu(1)
meansep = box / round(n ** (1 / 3.))
par = H.default_params(box, nmesh, n, TreeUseBH=1, TreeRcut=float(d["rcut"]))
with H.quiet_stdout():
    H.r_load(pos, mass, box, np.arange(1, n + 1, dtype=np.uint64))
    R.ref_sort_particles()
    _, _, sid = H.r_sorted_particles()
    H.r_tree(0.9)
    a0, _ = H.r_walk(par, meansep)
    pot0 = np.zeros(n); ft0 = np.zeros((n, 3))
    R.ref_get_particles(None, None, None, H.ptr(ft0), H.ptr(pot0))
    a1, _ = H.r_walk(dict(par, TreeUseBH=0), meansep, fulltree=d["oldacc_sorted"], gravpm=np.zeros((n, 3)))
np.savez(out, sorted_ids=sid, acc_bh=a0, acc_rel=a1, pot_bh=pot0, fulltree_bh=ft0)
