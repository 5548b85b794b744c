#!/usr/bin/env python
"""Regenerates tests/golden/pm_ref.npz from the reference's OWN PM code (gravpm.cpp + petapm.cpp + powerspectrum.cpp,
compiled unmodified into oracle/_ref/libshenqi_ref.so with only FFTW / heFFTe replaced by oracle/ref_shim/heffte.h; run
in the build container only).  Contents: a seeded 4096-particle clustered set of unequal masses in a box of 8, and
what gravpm_force (gravity.h:86) leaves in Part[].GravPM and adds to Part[].Potential for Nmesh 48, Asmth 1.5, plus
the rows of the power-spectrum file it writes (k, P, Nmodes, P/D1^2; powerspectrum.cpp:90-118)."""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H

assert H.has_reference_pm(), "build oracle/_ref first (make -C oracle)"
box, n, nmesh, asmth, G = 8.0, 4096, 48, 1.5, 43007.1
pos = H.random_clustered(n, box, 2024)
mass = (1.0 + 0.25 * (np.arange(n) % 3)).astype(np.float32)
pot, gravpm, rows = H.r_gravpm_force(pos, mass, box, nmesh, asmth, G)
np.savez_compressed(os.path.join(HERE, "pm_ref.npz"), pos=pos, mass=mass, box=np.float64(box), nmesh=np.int64(nmesh), asmth=np.float64(asmth),
                    G=np.float64(G), gravpm=gravpm, potential=pot, powerspectrum_rows=rows)
print("wrote pm_ref.npz: max |GravPM| %.6g" % np.abs(gravpm).max())
