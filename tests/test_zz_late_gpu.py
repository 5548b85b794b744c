"""GPU tests added after the round's GPU budget was spent: they have run only against a CPU stand-in context, so they live in
a file pytest collects LAST - under `pytest -x` a failure here hides no other test.
  * a 3 M-particle, tie-heavy radix sort (ADVICE: stability across lanes, rounds, warps and tiles);
  * the GPU PM against golden outputs of the reference's OWN gravpm_force (tests/golden/pm_ref.npz)."""
import numpy as np
import pytest
from helpers import o_keys, o_sort

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import shenqi_b200 as sq
    c = sq.Context(0)
    yield c
    c.close()


def make_params(box, nmesh, n, **kw):
    import shenqi_b200 as sq
    soft = kw.pop("GravitySoftening", (box / round(n ** (1 / 3.))) / 30.)
    return sq.Params.defaults(box, nmesh, soft, **kw)


def test_sort_stable_with_many_ties_large(ctx):
    """3 M particles on 125 distinct positions: in every radix pass long runs of equal digits are ranked by different
    lanes, rounds, warps and tiles; the permutation must be the stable one (ties stay in upload order), as the oracle's."""
    n = 3_000_000
    rng = np.random.default_rng(17)
    pos = np.floor(rng.random((n, 3)) * 5) * (8.0 / 5) + 0.01
    ctx.set_params(make_params(8.0, 32, n))
    ctx.upload_particles(pos, np.ones(n, np.float32))
    ctx.ph_sort()
    sk, perm = ctx.sorted_keys_perm()
    k = o_keys(pos, 8.0)
    operm = o_sort(k)
    assert np.array_equal(sk, k[operm])
    assert np.array_equal(perm.astype(np.int64), operm)


def test_pm_force_against_reference_gravpm_force(ctx):
    """the GPU PM against gravpm_force of the reference ITSELF: tests/golden/pm_ref.npz holds what the reference's own
    gravpm.cpp + petapm.cpp (compiled unmodified into oracle/_ref, only FFTW / heFFTe replaced by a stand-in header;
    tests/golden/make_golden_pm.py, tests/test_reference_pm_cpu.py) leave in GravPM and add to Potential for 4096 clustered
    particles of unequal mass at Nmesh 48, in input order."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "pm_ref.npz"))
    pos, mass, box, nmesh = g["pos"], g["mass"], float(g["box"]), int(g["nmesh"])
    P = make_params(box, nmesh, len(pos), Asmth=float(g["asmth"]), G=float(g["G"]))
    ctx.set_params(P)
    ctx.upload_particles(pos, mass); ctx.ph_sort()
    gpm, pot = ctx.gravpm_force()
    assert np.abs(gpm - g["gravpm"]).max() < 1e-10 * np.abs(g["gravpm"]).max()
    assert np.abs(pot - g["potential"]).max() < 1e-10 * np.abs(g["potential"]).max()
