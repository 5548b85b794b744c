#!/usr/bin/env python
"""MEASUREMENT AID (not product, not a test): times the REFERENCE's own CUDA tree walk - libgadget/gravshortcuda.cu +
treewalk2.cuh (thread-per-particle depth-first walk in FP64 over managed memory), compiled unmodified for sm_100a into
tests/dropin/_build/libshenqi_refgpu.so - on the bench ICs on this B200, next to the library's walk on the same
particles.  SURVEY 2b makes that kernel the bar.  Prints one JSON line per size."""
import argparse, ctypes, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench, helpers as H

ap = argparse.ArgumentParser()
ap.add_argument("--ngrid", type=int, nargs="+", default=[64, 128])
a = ap.parse_args()
H.REF_SO = os.path.join(ROOT, "tests", "dropin", "_build", "libshenqi_refgpu.so")
R = H.ref()
assert R is not None, "tests/dropin/_build/libshenqi_refgpu.so missing (make -C tests/dropin)"
import shenqi_b200 as sq
for ng in a.ngrid:
    pos, mass, par = bench.workload(ng, 2 * ng)
    n = len(pos); box = par["BoxSize"]; meansep = box / ng
    opar = H.default_params(box, 2 * ng, n, G=par["G"], TreeUseBH=1, TreeRcut=par["TreeRcut"], GravitySoftening=par["GravitySoftening"], rho0=par["rho0"])
    # ours, for |a_old| and the comparison
    ctx = sq.Context(0)
    P = sq.Params.defaults(box, 2 * ng, par["GravitySoftening"], G=par["G"], rho0=par["rho0"], TreeRcut=par["TreeRcut"], TreeUseBH=1)
    ctx.set_params(P); ctx.upload_particles(pos, mass); ctx.ph_sort(); ctx.tree_build()
    a0, _, st0 = ctx.grav_short_tree()
    P.TreeUseBH = 0; ctx.set_params(P); ctx.set_old_accel(a0, None)
    a1, _, st1 = ctx.grav_short_tree()
    ours_ms = st1.kernel_ms
    ctx.close()
    out = {"ngrid": ng, "particles": n, "ours_walk_ms": ours_ms, "interactions": int(st1.interactions)}
    with H.quiet_stdout():
        H.r_load(pos, mass, box, np.arange(1, n + 1, dtype=np.uint64))
        R.ref_sort_particles()
        _, _, sid = H.r_sorted_particles()
        inv = (sid - 1).astype(np.int64)
        H.r_tree(0.9)
        full = np.ascontiguousarray(a0[inv])
        for use_gpu, tag in ((1, "reference_gpu"), (0, "reference_cpu")):
            R.ref_set_use_gpu(use_gpu)
            times = []
            for rep in range(3 if use_gpu else 1):
                R.ref_set_accels(H.ptr(full), H.ptr(np.zeros((n, 3))))
                secs = R.ref_grav_short_tree(*H._r_args(dict(opar, TreeUseBH=0), meansep), H.c_dbl(opar["rho0"]), None)
                times.append(secs)
            acc = np.zeros((n, 3))
            R.ref_get_particles(None, None, None, H.ptr(acc), None)
            e = H.rel_err(a1[inv], acc)
            out[tag + "_walk_ms"] = [1e3 * t for t in times]
            out[tag + "_vs_ours_median_rel_err"] = float(np.median(e))
        out["cpu_threads"] = R.ref_omp_threads()
        R.ref_tree_free()
    out["reference_gpu_over_ours"] = min(out["reference_gpu_walk_ms"]) / ours_ms
    print(json.dumps(out), flush=True)
