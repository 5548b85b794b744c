#!/usr/bin/env python
"""Regenerates tests/golden/step_ref.npz from the reference's OWN drift_all_particles (libgadget/drift.cpp:81-99,
compiled unmodified into oracle/_ref/libshenqi_ref.so; run in the build container only).  Contents: a seeded
4096-particle set with velocities large enough that some coordinates wrap once or several times in either
direction, IsGarbage / Swallowed flags on a few particles, a random shift, and the reference's positions after the
drift.  build_active_sublist (libgadget/timestep.cpp:1353-1397) sits in a translation unit that does not link
stand-alone (cooling, star formation, black holes); its golden data is the literal std::copy_if semantics stated in
numpy below (stable selection by SubActivePredicate)."""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H


def inputs():
    rng = np.random.default_rng(20261104)
    n, box = 4096, 25.0
    pos = rng.random((n, 3)) * box
    pos[:8] = [[box, box, box], [1e-300, box, 0.5 * box], [box * (1 - 2 ** -52), 1e-9, box], [0.25 * box] * 3,
               [box, 1e-12, 1e-12], [0.5 * box, box, box], [box] * 3, [3.0, 4.0, 5.0]]
    vel = rng.standard_normal((n, 3)) * 40.0
    vel[rng.random(n) < 0.02] *= 50.0          # several box lengths in one drift
    vel[3] = 0.0
    flags = np.zeros(n, np.uint8)
    flags[rng.random(n) < 0.03] = 1            # IsGarbage
    flags[rng.random(n) < 0.03] |= 2           # Swallowed
    shift = np.array([0.37, -1.91, 12.5])
    return pos, vel, flags, 0.0173, shift, box


if __name__ == "__main__":
    pos, vel, flags, ddrift, shift, box = inputs()
    assert H.ref() is not None, "build oracle/_ref first (make -C oracle)"
    out = H.r_drift(pos, vel, flags, ddrift, shift, box)
    out0 = H.r_drift(pos, vel, None, -ddrift, np.zeros(3), box)
    np.savez_compressed(os.path.join(HERE, "step_ref.npz"), pos=pos, vel=vel, flags=flags, ddrift=np.float64(ddrift), shift=shift,
                        box=np.float64(box), pos_after=out, pos_after_back=out0)
    print("wrote step_ref.npz: %d particles, %d coordinates wrapped" % (len(pos), int((np.abs(out - pos - vel * ddrift - shift) > 1).sum())))
