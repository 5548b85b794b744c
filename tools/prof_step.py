#!/usr/bin/env python
"""Short driver for ncu: bootstrap |a_old| (Barnes-Hut pass + PM), then `--steps` device-resident steps of the hot
path on the bench workload (Zel'dovich ngrid^3, Nmesh = 2 ngrid).  Prints the stage timings of the last step."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import shenqi_b200 as sq

ap = argparse.ArgumentParser()
ap.add_argument("--ngrid", type=int, default=128)
ap.add_argument("--steps", type=int, default=1)
ap.add_argument("--group-width", type=int, default=32)
ap.add_argument("--no-pm", action="store_true")
ap.add_argument("--active-fraction", type=float, default=0.0, help="after the steps: tree + walk of a random active subset (hierarchical sub-step)")
ap.add_argument("--clustered", type=float, default=0.0, help="fraction of the particles in Hernquist halos (shenqi_b200.ics.clustered)")
a = ap.parse_args()
pos, mass, par = bench.workload(a.ngrid, 2 * a.ngrid)
if a.clustered > 0:
    from shenqi_b200.ics import clustered
    pos, mass, _ = clustered(a.ngrid, par["BoxSize"], a.clustered)
ctx = sq.Context(0)
ctx.set_group_width(a.group_width)
P = sq.Params.defaults(par["BoxSize"], 2 * a.ngrid, par["GravitySoftening"], G=par["G"], rho0=par["rho0"], TreeRcut=par["TreeRcut"], TreeUseBH=1)
ctx.set_params(P)
ctx.upload_particles(pos, mass)
ctx.ph_sort(); ctx.tree_build()
a0, _, _ = ctx.grav_short_tree()
g0, _ = ctx.gravpm_force()
P.TreeUseBH = 0
ctx.set_params(P)
ctx.set_old_accel(a0, g0)
for _ in range(a.steps):
    ctx.ph_sort(); ctx.tree_build()
    st = ctx.grav_short_tree_device()
    if not a.no_pm:
        ctx.gravpm_force_device()
if a.active_fraction > 0:
    # hierarchical stepping (timestep.cpp:291-300, 1375-1397): a tree of ONLY the active particles, walked by them
    import numpy as np, time
    rng = np.random.default_rng(7)
    act = np.flatnonzero(rng.random(len(pos)) < a.active_fraction).astype(np.int32)
    for _ in range(3):
        t0 = time.perf_counter()
        ctx.tree_build(act)
        tb = time.perf_counter()
        st = ctx.grav_short_tree_device(act)
        ctx.L.sq_sync(ctx._ctx)
        t1 = time.perf_counter()
    print(json.dumps({"active": int(len(act)), "fraction": a.active_fraction, "build_wall_ms": 1e3 * (tb - t0), "walk_wall_ms": 1e3 * (t1 - tb), "walk_ms": st.kernel_ms,
                      "build_kernels_ms": ctx.timings().tree_build_ms + ctx.timings().tree_moments_ms,
                      "interactions": st.interactions, "nodes": int(ctx.tree_info.numnodes)}))
tm = ctx.timings()
print(json.dumps({"n": len(pos), "interactions": st.interactions, "list_entries": st.list_entries, "nodes_tested": st.nodes_tested, "pair_tests": st.pair_tests, "exact_tests": st.exact_tests,
                  "ngroups": st.ngroups, "walk_ms": st.kernel_ms, "stages": {f: getattr(tm, f) for f, _ in tm._fields_}}))
ctx.close()
