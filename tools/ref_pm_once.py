#!/usr/bin/env python
"""TEST / BENCH INFRASTRUCTURE: runs gravpm_force of the reference's own PM build (oracle/_ref: gravpm.cpp + petapm.cpp
unmodified, stand-in FFT) once, in a process of its own, so that an abort inside the reference (endrun -> MPI_Abort)
cannot take the benchmark down.  Usage: ref_pm_once.py in.npz out.npz; in: pos, mass, box, nmesh, asmth, G; out: gravpm,
potential (input order), seconds (gravpm_force alone), threads."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H


def main():
    src, dst = sys.argv[1], sys.argv[2]
    g = np.load(src)
    pos, mass = np.ascontiguousarray(g["pos"], np.float64), np.ascontiguousarray(g["mass"], np.float32)
    n = len(pos)
    assert H.has_reference_pm(), "oracle/_ref was built without the reference's PM"
    R = H.ref()
    R.ref_set_threads(int(os.environ.get("OMP_NUM_THREADS", "0")) or (os.cpu_count() or 1))
    os.makedirs(H.REF_SCRATCH, exist_ok=True)
    with H.quiet_stdout():
        H.r_load(pos, mass, float(g["box"]))
        R.ref_set_potential(H.ptr(np.zeros(n)))
        R.ref_set_power_dir(H.REF_SCRATCH.encode())
        t0 = time.perf_counter()
        R.ref_gravpm_force(H.c_int(int(g["nmesh"])), H.c_dbl(float(g["asmth"])), H.c_dbl(float(g["G"])))
        secs = time.perf_counter() - t0
    gp = np.zeros((n, 3)); pot = np.zeros(n)
    R.ref_get_gravpm(H.ptr(gp))
    R.ref_get_particles(None, None, None, None, H.ptr(pot))
    np.savez(dst, gravpm=gp, potential=pot, seconds=np.float64(secs), threads=np.int64(R.ref_omp_threads()))


if __name__ == "__main__":
    main()
