"""BASELINE.json configurations under the reference's own code (oracle/_ref) inside `pytest -m gpu`.

The particle sets are bench.workload()'s (real box in kpc/h, real softening, Nmesh as configured):
  C1   dm-small size: Zel'dovich 64^3, Nmesh 192 (the default Nmesh rule of run.cpp:225-226 gives 3 x Ngrid)
  mid  128^3, Nmesh 256 (the bench ICs at a size the 16 host cores walk in seconds)
  C2   256^3, Nmesh 512 (the config BASELINE.json's metric is quoted on): everything that is O(N) is compared at full
       size (keys memcmp, sorted key sequence), the walk on a fixed 1/64 sample of the targets (every 64th warp-sized
       run of the PH order = 262 144 targets) against the FULL 256^3 tree of the reference, per-particle interaction
       counts included.
Procedure = SURVEY.md 8c: same ICs, same gravshort_tree_params, same GravPM / FullTreeGravAccel inputs for OldAcc, two
calls (Barnes-Hut, then the relative criterion, libgadget/tests/test_gravity.cpp:237-240).  Checks: keys memcmp; sorted
key sequence; canonical tree topology + moments 1e-12; interaction counts equal (per particle where both sides have
them); accelerations median <= 1e-5 and p99.9 <= 1e-4; error distribution against grav_short_pair
(gravshort-pair.c:22-127) the same for both.  TEST CODE: the reference build is the checker, never the product."""
import os
import sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from helpers import (ref, ptr, c_i64, c_dbl, c_int, quiet_stdout, canonical_tree, rel_err, default_params, r_load, r_tree, r_walk,
                     r_counts, r_pair, r_sorted_particles, _r_args)

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(ref() is None, reason="reference build (oracle/_ref) not present")]


def _setup(ng, nmesh):
    import bench
    import shenqi_b200 as sq
    pos, mass, par = bench.workload(ng, nmesh)
    P = sq.Params.defaults(par["BoxSize"], nmesh, par["GravitySoftening"], G=par["G"], rho0=par["rho0"], TreeRcut=par["TreeRcut"], TreeUseBH=1)
    opar = default_params(par["BoxSize"], nmesh, len(pos), G=par["G"], TreeUseBH=1, TreeRcut=par["TreeRcut"], GravitySoftening=par["GravitySoftening"], rho0=par["rho0"])
    return sq, pos, mass, par, P, opar


def reachable_nodes(nodes, firstnode):
    """number of nodes reachable from the root (the reference's numnodes also counts unused thread-cache slots)."""
    frontier = np.array([0], np.int64)
    tot = 0
    while len(frontier):
        tot += len(frontier)
        nd = nodes[frontier]
        inner = nd[((nd["flags"] >> 3) & 3) == 1]
        ch = inner["suns"].reshape(-1)
        frontier = ch[ch >= 0].astype(np.int64) - firstnode
    return tot


def _assert_parity(a, r, what):
    e = rel_err(a, r)
    med, p999 = float(np.median(e)), float(np.percentile(e, 99.9))
    assert med <= 1e-5 and p999 <= 1e-4, (what, med, p999)
    return med, p999


@pytest.mark.parametrize("ng,nmesh", [(64, 192), (128, 256)])
def test_baseline_config_against_reference_build(ng, nmesh):
    sq, pos, mass, par, P, opar = _setup(ng, nmesh)
    n, box, meansep = len(pos), par["BoxSize"], par["BoxSize"] / ng
    ids = np.arange(1, n + 1, dtype=np.uint64)
    c = sq.Context(0)
    try:
        c.set_params(P)
        c.upload_particles(pos, mass); c.ph_sort()
        keys = c.keys()
        skeys, perm = c.sorted_keys_perm()
        info = c.tree_build()
        nodes, first = c.tree_nodes(firstnode=n, upload_order=True)
        a0, _, st0 = c.grav_short_tree()
        g0, _ = c.gravpm_force()
        P.TreeUseBH = 0
        c.set_params(P)
        c.set_old_accel(a0, g0)
        a1, p1, st1 = c.grav_short_tree()
    finally:
        c.close()
    R = ref()
    with quiet_stdout():
        # 1. keys: memcmp
        rkeys = np.zeros(n, np.uint64)
        R.ref_peano_keys(ptr(pos), c_i64(n), c_dbl(box), ptr(rkeys))
        r_load(pos, mass, box, ids=ids)
        R.ref_sort_particles()
        _, _, rids = r_sorted_particles()
        inv = (rids - 1).astype(np.int64)
        rnodes, rfirst, _ = r_tree(0.9)
        r0, _ = r_walk(opar, meansep)
        c0, t0 = r_counts(opar, meansep)
        full = np.ascontiguousarray(a0[inv]); gpm = np.ascontiguousarray(g0[inv])
        opar0 = dict(opar, TreeUseBH=0)
        R.ref_set_accels(ptr(full), ptr(gpm))
        c1, t1 = r_counts(opar0, meansep)          # counts first: the walk overwrites FullTreeGravAccel on a full tree
        r1, _ = r_walk(opar0, meansep, fulltree=full, gravpm=gpm)
        pair, _ = r_pair(opar0, meansep)
        R.ref_tree_free()
    assert np.array_equal(keys, rkeys)
    # 2. sort order: identical key sequence (ties may permute)
    assert np.array_equal(skeys, rkeys[inv])
    # 3. topology: canonical cells + leaf membership by ID, moments 1e-12
    cells, parts, moms = canonical_tree(nodes, first, ids, box)
    rcells, rparts, rmoms = canonical_tree(rnodes, rfirst, rids, box)
    assert info.numnodes == len(rcells)          # reachable nodes (the reference's numnodes also counts unused cache slots)
    assert np.array_equal(cells, rcells) and np.array_equal(parts, rparts)
    assert (np.abs(moms - rmoms) / np.abs(rmoms).max(axis=0)).max() < 1e-12
    # 4. interaction counts: the reference's exactly, both criteria
    assert st0.interactions == int(t0), (st0.interactions, int(t0))
    assert st1.interactions == int(t1), (st1.interactions, int(t1))
    # 5. accelerations
    _assert_parity(a0[inv], r0, "Barnes-Hut")
    _assert_parity(a1[inv], r1, "relative")
    # 6. same error distribution against the grav_short_pair direct sum (metric of runtests.cpp:126-170)
    e_ours, e_ref = rel_err(a1[inv], pair), rel_err(r1, pair)
    assert abs(e_ours.mean() - e_ref.mean()) <= 1e-3 * e_ref.mean() + 1e-7
    assert abs(e_ours.max() - e_ref.max()) <= 1e-2 * e_ref.max() + 1e-6
    grid = np.sort(np.concatenate([e_ours, e_ref]))[:: max(1, 2 * n // 4096)]
    ks = np.abs(np.searchsorted(np.sort(e_ours), grid) - np.searchsorted(np.sort(e_ref), grid)).max() / n
    assert ks <= 2e-3, ks


def test_config2_full_size_sample_against_reference_build():
    """C2 at FULL size: 256^3, Nmesh 512.  Keys and the sorted key sequence of all 16.8 M particles against the reference;
    a fixed 1/64 sample of the targets (every 64th run of 32 consecutive PH-sorted particles) walked against the full
    tree by both, per-particle interaction counts and accelerations compared."""
    sq, pos, mass, par, P, opar = _setup(256, 512)
    n, box, meansep = len(pos), par["BoxSize"], par["BoxSize"] / 256
    c = sq.Context(0)
    try:
        c.set_params(P)
        c.upload_particles(pos, mass); c.ph_sort()
        keys = c.keys()
        skeys, perm = c.sorted_keys_perm()
        info = c.tree_build()
        ngroups = n // 32
        sel = (np.arange(0, ngroups, 64)[:, None] * 32 + np.arange(32)[None, :]).reshape(-1)
        act = np.ascontiguousarray(perm[sel].astype(np.int32))            # upload indices of the sampled targets
        a0 = np.zeros((n, 3)); a0, _, st0 = c.grav_short_tree(active=act, accel=a0)
        g0, _ = c.gravpm_force()
        P.TreeUseBH = 0
        c.set_params(P)
        c.set_old_accel(a0, g0)                                             # only the sampled rows matter
        a1 = np.zeros((n, 3)); a1, _, st1 = c.grav_short_tree(active=act, accel=a1)
    finally:
        c.close()
    assert st0.ntargets == len(act) == n // 64
    R = ref()
    ids = np.arange(1, n + 1, dtype=np.uint64)
    with quiet_stdout():
        rkeys = np.zeros(n, np.uint64)
        R.ref_peano_keys(ptr(pos), c_i64(n), c_dbl(box), ptr(rkeys))
        r_load(pos, mass, box, ids=ids)
        R.ref_sort_particles()
        _, _, rids = r_sorted_particles()
        inv = (rids - 1).astype(np.int64)                                  # reference sorted position -> upload index
        rnodes, rfirst, _ = r_tree(0.9)
        rnn = reachable_nodes(rnodes, rfirst)
        del rnodes
        where = np.zeros(n, np.int64); where[inv] = np.arange(n)           # upload index -> reference sorted position
        ract = np.ascontiguousarray(where[act].astype(np.int32))
        racc0 = np.zeros((n, 3)); cnt0 = np.zeros(len(ract), np.int64)
        R.ref_grav_short_tree_active.restype = c_dbl
        R.ref_grav_count_interactions_active.restype = c_i64
        t0 = R.ref_grav_count_interactions_active(*_r_args(opar, meansep), ptr(ract), c_i64(len(ract)), ptr(cnt0))
        R.ref_grav_short_tree_active(*_r_args(opar, meansep), c_dbl(opar["rho0"]), ptr(ract), c_i64(len(ract)), ptr(racc0))
        full = np.ascontiguousarray(a0[inv]); gpm = np.ascontiguousarray(g0[inv])
        R.ref_set_accels(ptr(full), ptr(gpm))
        opar0 = dict(opar, TreeUseBH=0)
        cnt1 = np.zeros(len(ract), np.int64)
        t1 = R.ref_grav_count_interactions_active(*_r_args(opar0, meansep), ptr(ract), c_i64(len(ract)), ptr(cnt1))
        racc1 = np.zeros((n, 3))
        R.ref_grav_short_tree_active(*_r_args(opar0, meansep), c_dbl(opar["rho0"]), ptr(ract), c_i64(len(ract)), ptr(racc1))
        R.ref_tree_free()
    assert np.array_equal(keys, rkeys)
    assert np.array_equal(skeys, rkeys[inv])
    assert info.numnodes == rnn, (info.numnodes, rnn)
    assert st0.interactions == int(t0) == int(cnt0.sum()), (st0.interactions, int(t0))
    assert st1.interactions == int(t1) == int(cnt1.sum()), (st1.interactions, int(t1))
    _assert_parity(a0[act], racc0[ract], "Barnes-Hut, 1/64 sample of 256^3")
    _assert_parity(a1[act], racc1[ract], "relative, 1/64 sample of 256^3")
