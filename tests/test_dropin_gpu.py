"""Drop-in test: the reference's own grav_short_tree(), compiled unmodified with USE_CUDA, calling the B200 plug-in
through the reference's GPU hook grav_short_tree_cuda() (shenqi_b200/host/gravshortcuda_b200.cpp), against the same
function on the reference's CPU path (oracle/_ref).  Reads like runtests.cpp:428-445 (CPU vs GPU tree, maxerr <= 1e-5
there; here the BASELINE.json bounds median <= 1e-5, p99.9 <= 1e-4)."""
import os, subprocess, sys
import numpy as np
import pytest
from helpers import ROOT, ref, random_clustered, rel_err, default_params, r_load, r_sorted_particles, r_tree, r_walk, ptr

pytestmark = pytest.mark.gpu
DROPIN = os.path.join(ROOT, "tests", "dropin", "_build", "libshenqi_dropin.so")


@pytest.mark.skipif(ref() is None or not os.path.exists(DROPIN), reason="reference build / drop-in harness did not travel")
def test_reference_grav_short_tree_with_b200_hook(tmp_path):
    box, n, nmesh, rcut = 8.0, 20000, 48, 7.0
    pos = random_clustered(n, box, 21)
    mass = (1 + np.random.default_rng(5).random(n)).astype(np.float32)
    meansep = box / round(n ** (1 / 3.))
    par = default_params(box, nmesh, n, TreeUseBH=1, TreeRcut=rcut)
    # CPU path of the reference
    R = ref()
    r_load(pos, mass, box, np.arange(1, n + 1, dtype=np.uint64))
    R.ref_sort_particles()
    _, _, sid = r_sorted_particles()
    r_tree(0.9)
    c0, _ = r_walk(par, meansep)
    cpot = np.zeros(n)
    R.ref_get_particles(None, None, None, None, ptr(cpot))
    c1, _ = r_walk(dict(par, TreeUseBH=0), meansep, fulltree=c0, gravpm=np.zeros((n, 3)))
    R.ref_tree_free()
    # the same calls with UseGPU=true in the drop-in build
    inp, out = str(tmp_path / "in.npz"), str(tmp_path / "out.npz")
    np.savez(inp, pos=pos, mass=mass, box=box, nmesh=nmesh, rcut=rcut, oldacc_sorted=c0)
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tests", "dropin", "run_dropin.py"), inp, out])
    g = np.load(out)
    assert np.array_equal(g["sorted_ids"], sid)
    for a, b in ((g["acc_bh"], c0), (g["acc_rel"], c1)):
        e = rel_err(a, b)
        assert np.median(e) <= 1e-5 and np.percentile(e, 99.9) <= 1e-4, (np.median(e), np.percentile(e, 99.9))
    assert np.array_equal(g["fulltree_bh"], g["acc_bh"])            # postprocess copies Accel into FullTreeGravAccel
    assert np.median(np.abs(g["pot_bh"] - cpot) / np.abs(cpot)) < 1e-5


DROPIN_TREE = os.path.join(ROOT, "tests", "dropin", "_build", "libshenqi_dropin_tree.so")


@pytest.mark.skipif(ref() is None or not os.path.exists(DROPIN_TREE), reason="reference build / drop-in harness did not travel")
@pytest.mark.parametrize("garbage", [False, True])
def test_reference_application_with_all_three_bindings(tmp_path, garbage):
    """SURVEY 8b through the reference's own symbols: force_tree_full (forcetree_b200.cpp) must hand back a ForceTree whose
    NODE[] the reference's OWN CPU grav_short_tree(UseGPU=false) walks to the reference's accelerations; gravpm_force
    (gravpm_b200.cpp) must fill Part[i].GravPM and ADD to Part[i].Potential (gravpm.cpp:489-500); grav_short_tree(UseGPU=true)
    then walks the cached device tree.  garbage: IsGarbage / swallowed black holes stay out of the tree (forcetree.cpp:805)."""
    from helpers import canonical_tree, o_pm_force, r_counts, quiet_stdout
    box, n, nmesh, rcut = 8.0, 20000, 48, 7.0
    pos = random_clustered(n, box, 33)
    mass = (1 + np.random.default_rng(6).random(n)).astype(np.float32)
    rng = np.random.default_rng(9)
    pot0 = rng.standard_normal(n)
    meansep = box / round(n ** (1 / 3.))
    par = default_params(box, nmesh, n, TreeUseBH=1, TreeRcut=rcut)
    ids = np.arange(1, n + 1, dtype=np.uint64)
    flags = np.zeros(n, np.uint8); types = np.ones(n, np.uint8)
    if garbage:
        g = rng.choice(n, 300, replace=False)
        flags[g[:200]] = 1                       # IsGarbage
        flags[g[200:]] = 2; types[g[200:]] = 5   # swallowed black holes
    keep = flags == 0
    # the reference, pure (oracle/_ref)
    R = ref()
    with quiet_stdout():
        r_load(pos, mass, box, ids)
        R.ref_sort_particles()
        _, _, sid = r_sorted_particles()
        inv = (sid - 1).astype(np.int64)
        R.ref_set_flags(ptr(np.ascontiguousarray(flags[inv])), ptr(np.ascontiguousarray(types[inv])))
        rnodes, rfirst, _ = r_tree(0.9)
        c0, _ = r_walk(par, meansep)
        rcnt, rtot = r_counts(par, meansep)
        R.ref_tree_free()
    inp, out = str(tmp_path / "in.npz"), str(tmp_path / "out.npz")
    np.savez(inp, pos=pos, mass=mass, box=box, nmesh=nmesh, rcut=rcut, pot0=pot0, flags=flags, types=types)
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tests", "dropin", "run_dropin_tree.py"), inp, out])
    g = np.load(out)
    assert np.array_equal(g["sorted_ids"], sid)
    # ---- tree: flags, counts, topology, links usable by the reference's walk
    assert int(g["tree_flags"]) == 7                       # moments_computed | full_particle_tree | Father allocated
    assert int(g["numparticles"]) == int(keep.sum())
    cells, parts, moms = canonical_tree(g["nodes"], int(g["firstnode"]), sid, box)
    rcells, rparts, rmoms = canonical_tree(rnodes, rfirst, sid, box)
    assert np.array_equal(cells, rcells) and np.array_equal(parts, rparts)
    assert (np.abs(moms - rmoms) / np.abs(rmoms).max(axis=0)).max() < 1e-12
    nd = g["nodes"]; fa = g["fathers"]
    ks = keep[inv]
    assert np.all(((nd["flags"][fa[ks]] >> 3) & 3) == 0)   # Father[i] is a particle node ...
    assert np.all((nd["suns"][fa[ks]] == np.flatnonzero(ks)[:, None]).any(axis=1))   # ... that holds particle i
    # ---- the reference's own CPU walk over the GPU-built image: the reference's interactions and accelerations
    assert int(g["total"]) == int(rtot)
    e = rel_err(g["acc_cpu_on_b200_tree"][ks], c0[ks])
    assert e.max() <= 1e-10, e.max()
    # ---- the hook on the cached device tree
    e = rel_err(g["acc_gpu"][ks], c0[ks])
    assert np.median(e) <= 1e-5 and np.percentile(e, 99.9) <= 1e-4, (np.median(e), np.percentile(e, 99.9))
    # ---- gravpm_force: GravPM filled, Potential added to
    spos = np.ascontiguousarray(pos[inv][ks]); smass = np.ascontiguousarray(mass[inv][ks])
    opot, oacc = o_pm_force(spos, smass, box, nmesh, 1.5, par["G"])
    assert np.abs(g["gravpm"][ks] - oacc).max() <= 1e-9 * np.abs(oacc).max()
    assert np.abs((g["pot_after_pm"][ks] - pot0[inv][ks]) - opot).max() <= 1e-9 * np.abs(opot).max()
    if garbage:
        assert np.all(g["gravpm"][~ks] == 0)
