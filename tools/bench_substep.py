#!/usr/bin/env python
"""Hierarchical sub-steps (BASELINE config 5's regime, libgadget/timestep.cpp:425-472) on the 256^3 bench load with
everything device-resident: per time bin b (active fraction 2^-k), wall time of
    sub-list (sq_active_sublist_device) -> tree of the sub-list (sq_tree_build) -> walk (sq_tree_walk_device)
    -> kick through the list (sq_kick_device)
with the list as a DEVICE pointer, next to the same sub-step driven by a host copy of the list.  One JSON line."""
import argparse, json, os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import shenqi_b200 as sq

ap = argparse.ArgumentParser()
ap.add_argument("--ngrid", type=int, default=256)
a = ap.parse_args()
pos, mass, par = bench.workload(a.ngrid, 2 * a.ngrid)
n = len(pos)
ctx = sq.Context(0)
P = sq.Params.defaults(par["BoxSize"], 2 * a.ngrid, par["GravitySoftening"], G=par["G"], rho0=par["rho0"], TreeRcut=par["TreeRcut"], TreeUseBH=1)
ctx.set_params(P)
ctx.upload_particles(pos, mass)
ctx.ph_sort(); ctx.tree_build()
a0, _, _ = ctx.grav_short_tree()
P.TreeUseBH = 0
ctx.set_params(P)
ctx.set_old_accel(a0, np.zeros_like(a0))
# bins with geometric occupation: bin 10 holds 1/2 of the particles, bin 9 1/4, ... bin 1 2^-10 (random in space)
rng = np.random.default_rng(7)
u = rng.random(n)
tbh = np.clip(10 + np.floor(np.log2(np.maximum(u, 2.0 ** -10))).astype(np.int64) + 1, 1, 10).astype(np.uint8)
timebin = torch.from_numpy(tbh).cuda()
acc = torch.from_numpy(a0).cuda()
vel = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
out = torch.empty(n, dtype=torch.int32, device="cuda")
torch.cuda.synchronize()
sync = lambda: ctx._check(ctx.L.sq_sync(ctx._ctx))
rows = []
for b in (10, 9, 7, 4, 1):
    for mode in ("device_list", "host_list"):
        best = None
        for rep in range(3):
            sync(); t0 = time.perf_counter()
            m = ctx.build_active_sublist_device(n, timebin.data_ptr(), b, 1 << b, out.data_ptr())
            t1 = time.perf_counter()
            if mode == "device_list":
                ctx.tree_build_device_list(out.data_ptr(), m); sync(); t2 = time.perf_counter()
                st = ctx.grav_short_tree_device_list(out.data_ptr(), m); sync(); t3 = time.perf_counter()
            else:
                lst = out[:m].cpu().numpy()
                ctx.tree_build(lst); sync(); t2 = time.perf_counter()
                st = ctx.grav_short_tree_device(lst); sync(); t3 = time.perf_counter()
            ctx.kick_device(vel.data_ptr(), acc.data_ptr(), m, 1e-3, active_ptr=out.data_ptr()); sync(); t4 = time.perf_counter()
            r = dict(bin=b, mode=mode, active=m, fraction=m / n, sublist_ms=1e3 * (t1 - t0), build_ms=1e3 * (t2 - t1), walk_wall_ms=1e3 * (t3 - t2),
                     walk_kernel_ms=st.kernel_ms, kick_ms=1e3 * (t4 - t3), total_ms=1e3 * (t4 - t0), interactions=st.interactions)
            if best is None or r["total_ms"] < best["total_ms"]:
                best = r
        rows.append(best)
print(json.dumps({"particles": n, "timing": "host wall clock around synchronised calls, best of 3", "substeps": rows}))
ctx.close()
