#!/usr/bin/env python
"""GPU box check: total interaction count and acceleration parity of the walk against the reference's own CPU code
(oracle/_ref) on the bench workload at a size the CPU finishes in seconds (TEST TOOL, not product)."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench, helpers as H
import shenqi_b200 as sq

ap = argparse.ArgumentParser()
ap.add_argument("--ngrid", type=int, default=64)
ap.add_argument("--with-pm", action="store_true", help="|a_old| from tree + PM (as bench.py / prof_step.py) instead of the tree alone")
ap.add_argument("--clustered", type=float, default=0.0, help="fraction of the particles in Hernquist halos (shenqi_b200.ics.clustered)")
a = ap.parse_args()
pos, mass, par = bench.workload(a.ngrid, 2 * a.ngrid)
if a.clustered > 0:
    from shenqi_b200.ics import clustered
    pos, mass, _ = clustered(a.ngrid, par["BoxSize"], a.clustered)
n = len(pos); box = par["BoxSize"]
ctx = sq.Context(0)
P = sq.Params.defaults(box, 2 * a.ngrid, par["GravitySoftening"], G=par["G"], rho0=par["rho0"], TreeRcut=par["TreeRcut"], TreeUseBH=1)
ctx.set_params(P); ctx.upload_particles(pos, mass); ctx.ph_sort(); ctx.tree_build()
a0, _, st0 = ctx.grav_short_tree()
g0 = ctx.gravpm_force()[0] if a.with_pm else None
P.TreeUseBH = 0; ctx.set_params(P); ctx.set_old_accel(a0, g0)
a1, _, st1 = ctx.grav_short_tree()
meansep = box / a.ngrid
opar = H.default_params(box, 2 * a.ngrid, n, G=par["G"], TreeUseBH=1, TreeRcut=par["TreeRcut"], GravitySoftening=par["GravitySoftening"], rho0=par["rho0"])
with H.quiet_stdout():
    H.r_load(pos, mass, box, ids=np.arange(1, n + 1, dtype=np.uint64)); H.ref().ref_sort_particles(); H.r_tree(0.9)
    _, _, ids = H.r_sorted_particles()
    r0, _ = H.r_walk(opar, meansep)
    _, t0 = H.r_counts(opar, meansep)
    inv = (ids - 1).astype(np.int64)
    full = np.zeros((n, 3)); full[:] = a0[inv]
    opar["TreeUseBH"] = 0
    # counts first: the walk itself overwrites FullTreeGravAccel (the |a_old| input) on a full tree
    gpm = np.zeros((n, 3))
    if g0 is not None:
        gpm[:] = g0[inv]
    H.ref().ref_set_accels(H.ptr(full), H.ptr(gpm))
    _, t1 = H.r_counts(opar, meansep)
    r1, _ = H.r_walk(opar, meansep, fulltree=full, gravpm=gpm)
e0 = H.rel_err(a0[inv], r0); e1 = H.rel_err(a1[inv], r1)
print(json.dumps({"n": n, "bh": {"gpu": st0.interactions, "ref": int(t0), "median": float(np.median(e0)), "p999": float(np.percentile(e0, 99.9))},
                  "rel": {"gpu": st1.interactions, "ref": int(t1), "median": float(np.median(e1)), "p999": float(np.percentile(e1, 99.9)),
                          "pair_tests": st1.pair_tests, "exact_tests": st1.exact_tests}}))
