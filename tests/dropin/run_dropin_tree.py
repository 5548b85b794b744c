#!/usr/bin/env python
"""Runs the reference application with ALL THREE B200 bindings linked (tests/dropin/_build/libshenqi_dropin_tree.so:
forcetree_b200.cpp, gravpm_b200.cpp, gravshortcuda_b200.cpp + the reference's own objects) on a particle set from an
.npz: gravpm_force, force_tree_full, then the reference's OWN CPU grav_short_tree on the NODE image the GPU build handed
back, then grav_short_tree(UseGPU=true) on the device tree the build left in the context.  Own process: the library
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
H.REF_SO = os.path.join(HERE, "_build", "libshenqi_dropin_tree.so")
R = H.ref()
assert R is not None, "tests/dropin/_build/libshenqi_dropin_tree.so missing (make -C tests/dropin)"
meansep = box / round(n ** (1 / 3.))
par = H.default_params(box, nmesh, n, TreeUseBH=1, TreeRcut=float(d["rcut"]))
res = {}
with H.quiet_stdout():
    H.r_load(pos, mass, box, np.arange(1, n + 1, dtype=np.uint64))
    R.ref_sort_particles()
    _, _, sid = H.r_sorted_particles()
    inv = (sid - 1).astype(np.int64)
    if "flags" in d.files:
        R.ref_set_flags(H.ptr(np.ascontiguousarray(d["flags"][inv])), H.ptr(np.ascontiguousarray(d["types"][inv])))
    R.ref_set_potential(H.ptr(np.ascontiguousarray(d["pot0"][inv])))
    # ---- gravpm_b200.cpp
    R.ref_gravpm_force(ctypes.c_int(nmesh), ctypes.c_double(1.5), ctypes.c_double(par["G"]))
    gpm = np.zeros((n, 3)); pot_pm = np.zeros(n)
    R.ref_get_gravpm(H.ptr(gpm)); R.ref_get_particles(None, None, None, None, H.ptr(pot_pm))
    # ---- forcetree_b200.cpp: force_tree_full -> the reference's NODE image
    nodes, first, _ = H.r_tree(0.9)
    fathers = np.zeros(n, np.int32); R.ref_tree_fathers(H.ptr(fathers))
    res.update(tree_flags=R.ref_tree_flags(), numparticles=R.ref_tree_numparticles2(), firstnode=first)
    # ---- the reference's own CPU walk over that image
    R.ref_set_use_gpu(0)
    a_cpu, _ = H.r_walk(par, meansep)
    cnt, tot = H.r_counts(par, meansep)
    # ---- grav_short_tree(UseGPU = true): the hook walks the device tree that force_tree_full left behind
    R.ref_set_use_gpu(1)
    a_gpu, _ = H.r_walk(par, meansep)
    pot_tree = np.zeros(n); R.ref_get_particles(None, None, None, None, H.ptr(pot_tree))
np.savez(out, sorted_ids=sid, gravpm=gpm, pot_after_pm=pot_pm, nodes=nodes, fathers=fathers, acc_cpu_on_b200_tree=a_cpu, counts=cnt, total=tot,
         acc_gpu=a_gpu, pot_tree=pot_tree, **res)
