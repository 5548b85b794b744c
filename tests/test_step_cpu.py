"""SURVEY 8(f2) on the CPU: the oracle's drift / kick / active sub-list restatement against the reference's own
drift_all_particles (oracle/_ref, when built), the committed golden outputs of that code (tests/golden/step_ref.npz,
made by tests/golden/make_golden_step.py), and the literal semantics of libgadget/timestep.cpp:132-139, 962-968,
1353-1397.  No GPU."""
import os
import numpy as np
import pytest
from helpers import ROOT, oracle, ref, o_drift, o_kick, o_active_sublist, r_drift, o_gravity_timebins, timebin_params, c_i64, c_int, c_dbl, ptr

needs_ref = pytest.mark.skipif(ref() is None, reason="oracle/_ref was not built (no /root/reference at build time)")
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "step_ref.npz"))


def gold_inputs():
    return GOLD["pos"], GOLD["vel"], GOLD["flags"], float(GOLD["ddrift"]), GOLD["shift"], float(GOLD["box"])


def test_oracle_drift_equals_golden_reference_output():
    """bit for bit: Pos += Vel * ddrift + shift with separately rounded product and sums, wrap into (0, Box] (drift.cpp:68-79)."""
    pos, vel, flags, ddrift, shift, box = gold_inputs()
    out, bad = o_drift(pos, vel, (flags != 0).astype(np.uint8), ddrift, shift, box)
    assert bad == 0
    assert np.array_equal(out, GOLD["pos_after"])
    assert out.min() > 0 and out.max() <= box
    skipped = flags != 0
    assert skipped.sum() > 50       # garbage / swallowed particles only follow the shift (drift.cpp:20-29)
    back, _ = o_drift(pos, vel, None, -ddrift, np.zeros(3), box)
    assert np.array_equal(back, GOLD["pos_after_back"])


@needs_ref
def test_oracle_drift_matches_reference_live():
    rng = np.random.default_rng(5)
    n, box = 20000, 250000.0
    pos = rng.random((n, 3)) * box
    vel = rng.standard_normal((n, 3)) * 3e5
    flags = (rng.random(n) < 0.05).astype(np.uint8) * rng.integers(1, 4, n).astype(np.uint8)
    for ddrift, shift in ((0.11, np.zeros(3)), (2.5, np.array([-7e4, 1e5, 3.0]))):
        exp = r_drift(pos, vel, flags, ddrift, shift, box)
        out, bad = o_drift(pos, vel, (flags != 0).astype(np.uint8), ddrift, shift, box)
        assert bad == 0 and np.array_equal(out, exp)


def test_oracle_drift_counts_nonfinite():
    pos = np.full((4, 3), 1.0); vel = np.zeros((4, 3)); vel[2, 1] = np.inf; vel[3, 0] = np.nan
    out, bad = o_drift(pos, vel, None, 1.0, None, 10.0)
    assert bad == 2 and np.array_equal(out[:2], pos[:2])


def test_oracle_kick():
    rng = np.random.default_rng(6)
    n = 5000
    vel = rng.standard_normal((n, 3)); acc = rng.standard_normal((n, 3)) * 1e3
    skip = (rng.random(n) < 0.1).astype(np.uint8)
    act = np.sort(rng.choice(n, n // 3, replace=False)).astype(np.int32)
    exp = vel.copy()
    sel = act[skip[act] == 0]
    exp[sel] = vel[sel] + acc[sel] * 0.0371
    assert np.array_equal(o_kick(vel, acc, act, skip, 0.0371), exp)
    exp = vel.copy(); exp[skip == 0] += acc[skip == 0] * -2.0
    assert np.array_equal(o_kick(vel, acc, None, skip, -2.0), exp)


def test_oracle_is_timebin_active():
    L = oracle()
    assert L.sqo_is_timebin_active(0, 12345) == 1 and L.sqo_is_timebin_active(20, 0) == 1 and L.sqo_is_timebin_active(-1, 7) == 1
    for b in range(1, 46):
        assert L.sqo_is_timebin_active(b, 1 << b) == 1
        assert L.sqo_is_timebin_active(b, (1 << b) + (1 << (b - 1))) == 0
        assert L.sqo_is_timebin_active(b, 3 << b) == 1
    assert L.sqo_is_timebin_active(45, 1 << 46) == 1


def numpy_sublist(active, n, timebin, skip, maxtimebin, ti):
    """std::copy_if over the active list with SubActivePredicate (timestep.cpp:1359-1369)."""
    cand = np.arange(n, dtype=np.int32) if active is None else active
    b = timebin[cand].astype(np.int64)
    ok = (b <= maxtimebin) & ((b <= 0) | (ti <= 0) | (ti % (np.int64(1) << np.maximum(b, 0)) == 0))
    if skip is not None:
        ok &= skip[cand] == 0
    return cand[ok]


@pytest.mark.parametrize("with_list", [False, True])
def test_oracle_active_sublist(with_list):
    rng = np.random.default_rng(7)
    n = 30000
    timebin = rng.integers(0, 12, n).astype(np.uint8)
    skip = (rng.random(n) < 0.05).astype(np.uint8)
    act = np.sort(rng.choice(n, n // 2, replace=False)).astype(np.int32) if with_list else None
    na = n if act is None else len(act)
    for ti in (0, 1 << 5, 3 << 4, 1 << 11, (1 << 11) + 64):
        for mtb in (0, 3, 6, 46):
            got = o_active_sublist(act, na, timebin, skip, mtb, ti)
            assert np.array_equal(got, numpy_sublist(act, n, timebin, skip, mtb, ti))
    assert len(o_active_sublist(act, 0, timebin, skip, 46, 0)) == 0


def test_oracle_round_down_power_of_two_known_answers():
    """libgadget/tests/test_timebinmgr.cpp:143-145, and get_timestep_bin (timestep.cpp:1236-1250)."""
    L = oracle()
    TB = 1 << 46
    assert L.sqo_round_down_power_of_two(TB) == TB
    assert L.sqo_round_down_power_of_two(TB + 1) == TB
    assert L.sqo_round_down_power_of_two(TB - 1) == TB // 2
    assert L.sqo_round_down_power_of_two(0) == 0 and L.sqo_round_down_power_of_two(1) == 1 and L.sqo_round_down_power_of_two(-5) == -4
    assert [L.sqo_get_timestep_bin(x) for x in (0, 1, 2, 3, 4, 1 << 20, (1 << 21) - 1, TB)] == [0, 0, 1, 1, 2, 20, 20, 46]


def numpy_timebins(P, accel, gravpm):
    """The same chain in numpy float64 (elementwise IEEE operations in source order)."""
    a2inv = 1 / (P.atime * P.atime)
    a = a2inv * accel + a2inv * gravpm
    ac2 = a[:, 0] * a[:, 0] + a[:, 1] * a[:, 1] + a[:, 2] * a[:, 2]
    ac2[ac2 == 0] = 1e-60
    dloga = np.sqrt(2 * P.ErrTolIntAccuracy * P.atime * P.softening / np.sqrt(ac2)) * P.hubble
    dloga = np.maximum(dloga, P.MinSizeTimestep)
    target = dloga + P.loga
    adv = (target >= P.sync_loga[1]) & bool(P.can_advance)
    s0 = np.where(adv, P.sync_loga[1], P.sync_loga[0]); s1 = np.where(adv, P.sync_loga[2], P.sync_loga[1])
    tnew = ((P.lastsnap + adv.astype(np.int64)) << 46).astype(np.float64) + (target - s0) / ((s1 - s0) / float(1 << 46))
    tip = np.where(np.abs(tnew) < 2.0 ** 63, tnew, -2.0 ** 63).astype(np.int64)
    dti = tip - P.Ti_Current
    dti[(dti > P.dti_max) | (dti < 0)] = P.dti_max
    dti = np.minimum(dti, 1 << 46)
    bins = np.where(dti <= 1, 0, np.floor(np.log2(np.maximum(dti, 1).astype(np.float64))).astype(np.int64))
    # log2 of an int64 near a power of two can round up: fix by the exact integer test
    bins = np.where((np.int64(1) << bins) > np.maximum(dti, 1), bins - 1, bins)
    return np.minimum(bins, P.largest_active)


def test_oracle_gravity_timebins():
    rng = np.random.default_rng(8)
    n = 40000
    acc = rng.standard_normal((n, 3)) * 10.0 ** rng.uniform(-2, 6, (n, 1))
    pm = rng.standard_normal((n, 3)) * 10.0
    acc[:3] = 0; pm[:3] = 0                                  # ac2 == 0 -> 1e-60 (timestep.cpp:1024)
    skip = (rng.random(n) < 0.05).astype(np.uint8)
    for kw in (dict(), dict(Ti_Current=(1 << 45) + (1 << 30), loga=np.log(0.1) + ((1 << 45) + (1 << 30)) * (np.log(0.111) - np.log(0.1)) / (1 << 46)),
               dict(largest_active=33), dict(dti_max=1 << 31), dict(MinSizeTimestep=1e-4), dict(can_advance=0)):
        P = timebin_params(**kw)
        tb, counts, nbad = o_gravity_timebins(P, acc, pm, None, skip, n)
        exp = numpy_timebins(P, acc, pm)
        sel = skip == 0
        assert np.array_equal(tb[sel], exp[sel]) and np.all(tb[~sel] == 255)
        assert np.array_equal(counts, np.bincount(exp[sel], minlength=47)) and counts.sum() == sel.sum()
        assert len(np.unique(tb[sel])) >= 3 or "largest_active" in kw or "dti_max" in kw
        assert nbad == 0
    act = np.sort(rng.choice(n, n // 4, replace=False)).astype(np.int32)
    P = timebin_params()
    tb, counts, _ = o_gravity_timebins(P, acc, pm, act, skip, n)
    exp = numpy_timebins(P, acc, pm)
    sel = act[skip[act] == 0]
    assert np.array_equal(tb[sel], exp[sel]) and counts.sum() == len(sel) and (tb != 255).sum() == len(sel)
    # dti_max == 0 -> dti 0 -> bin 0, reported as bad (timestep.cpp:161-162, 369-370)
    tb, counts, nbad = o_gravity_timebins(timebin_params(dti_max=0), acc, pm, None, None, n)
    assert nbad == n and counts[0] == n


def test_pushdown_rule_host_equals_oracle():
    """timestep.cpp:379-398: product host helper (shenqi_b200.pushdown_bin) against the C restatement."""
    import shenqi_b200 as sq
    rng = np.random.default_rng(9)
    L = oracle()
    for trial in range(200):
        counts = np.zeros(47, np.int64)
        top = int(rng.integers(2, 30))
        occ = rng.integers(0, 2, top + 1).astype(bool)
        counts[:top + 1] = rng.integers(0, 10 ** int(rng.integers(1, 7)), top + 1) * occ
        for isPM in (0, 1):
            c2 = counts.copy()
            exp = L.sqo_pushdown_bin(ptr(c2), c_int(40), c_int(isPM))
            got, c3 = sq.pushdown_bin(counts.tolist(), 40, isPM)
            assert got == exp and c3 == c2.tolist()


# ------------------------------------------------------------------------------------------ pins to the reference's own timestep.cpp
needs_ref = pytest.mark.skipif(ref() is None, reason="reference build (oracle/_ref) not present")


def _ref_load_for_timestep(n, gravpm, flags, seed=3):
    """n particles in the reference's particle table with the given GravPM and IsGarbage / Swallowed flags."""
    from helpers import r_load, quiet_stdout
    rng = np.random.default_rng(seed)
    pos = rng.random((n, 3)) * 8.0
    with quiet_stdout():
        r_load(pos, np.ones(n, np.float32), 8.0)
        ref().ref_set_accels(None, ptr(np.ascontiguousarray(gravpm)))
        ref().ref_set_flags(ptr(np.ascontiguousarray(flags)), None)


@needs_ref
def test_timebins_equal_reference_timestep_cpp():
    """The bin loop of hierarchical_gravity_accelerations (timestep.cpp:351-377) run with the reference's OWN static
    functions (oracle/ref_timestep.cpp includes timestep.cpp where it lies): get_timestep_gravity_dloga,
    convert_timestep_to_ti -> TimeBinMgr::dti_from_dloga, round_down_power_of_two, get_timestep_bin.  The restatement
    (and with it the GPU kernel, tests/test_step_gpu.py) must give the same bin for every particle and the same counts."""
    from helpers import quiet_stdout
    import math
    R = ref()
    rng = np.random.default_rng(8)
    n = 40000
    acc = rng.standard_normal((n, 3)) * 10.0 ** rng.uniform(-2, 6, (n, 1))
    pm = rng.standard_normal((n, 3)) * 10.0
    acc[:3] = 0; pm[:3] = 0
    flags = np.where(rng.random(n) < 0.05, rng.integers(1, 3, n), 0).astype(np.uint8)      # IsGarbage (1) or Swallowed (2)
    skip = (flags != 0).astype(np.uint8)
    _ref_load_for_timestep(n, pm, flags)
    sync = np.array([math.log(0.1), math.log(0.111), math.log(0.125)])
    R.ref_ts_gravity_timebins.restype = c_i64
    R.ref_ts_loga_from_ti.restype = c_dbl
    soft = 1.0 / 30
    for kw in (dict(), dict(Ti_Current=(1 << 45) + (1 << 30)), dict(largest_active=33), dict(dti_max=1 << 31), dict(MinSizeTimestep=1e-4),
               dict(nsync=2), dict(dti_max=0)):
        nsync = kw.pop("nsync", 3)
        Ti = kw.get("Ti_Current", 0)
        with quiet_stdout():
            R.ref_ts_setup(c_dbl(0.02), c_dbl(kw.get("MinSizeTimestep", 0.0)), ptr(sync), c_int(nsync))
            R.ref_ts_set_softening(c_dbl(soft))
        loga = R.ref_ts_loga_from_ti(c_i64(Ti))
        P = timebin_params(loga=loga, can_advance=1 if nsync == 3 else 0, softening=soft, **kw)
        rtb = np.zeros(n, np.uint8); rcounts = np.zeros(47, np.int64)
        with quiet_stdout():
            R.ref_ts_set_timebins(ptr(np.full(n, 255, np.uint8)))
            rbad = R.ref_ts_gravity_timebins(ptr(acc), None, c_i64(n), c_dbl(P.atime), c_dbl(P.hubble), c_i64(P.Ti_Current), c_i64(P.dti_max),
                                             c_int(P.largest_active), ptr(rtb), ptr(rcounts))
        tb, counts, nbad = o_gravity_timebins(P, acc, pm, None, skip, n)
        assert np.array_equal(tb, rtb), kw
        assert np.array_equal(counts, rcounts) and nbad == rbad, kw
    # an active list
    act = np.sort(rng.choice(n, n // 4, replace=False)).astype(np.int32)
    with quiet_stdout():
        R.ref_ts_setup(c_dbl(0.02), c_dbl(0.0), ptr(sync), c_int(3))
        R.ref_ts_set_timebins(ptr(np.full(n, 255, np.uint8)))
        P = timebin_params(loga=R.ref_ts_loga_from_ti(c_i64(0)), softening=soft)
        rtb = np.zeros(n, np.uint8); rcounts = np.zeros(47, np.int64)
        R.ref_ts_gravity_timebins(ptr(acc), ptr(act), c_i64(len(act)), c_dbl(P.atime), c_dbl(P.hubble), c_i64(0), c_i64(P.dti_max), c_int(P.largest_active), ptr(rtb), ptr(rcounts))
    tb, counts, _ = o_gravity_timebins(P, acc, pm, act, skip, n)
    assert np.array_equal(tb, rtb) and np.array_equal(counts, rcounts)


@needs_ref
@pytest.mark.parametrize("with_list", [False, True])
def test_active_sublist_and_kick_equal_reference_timestep_cpp(with_list):
    """build_active_sublist (timestep.cpp:1375-1397, std::copy_if with SubActivePredicate) and the kick loop of
    apply_hierarchical_grav_kick (:270-279, do_grav_short_range_kick :962-968), the reference's own code."""
    from helpers import quiet_stdout
    R = ref()
    rng = np.random.default_rng(7)
    n = 30000
    timebin = rng.integers(0, 12, n).astype(np.uint8)
    flags = np.where(rng.random(n) < 0.05, rng.integers(1, 3, n), 0).astype(np.uint8)
    skip = (flags != 0).astype(np.uint8)
    _ref_load_for_timestep(n, np.zeros((n, 3)), flags)
    R.ref_ts_active_sublist.restype = c_i64
    with quiet_stdout():
        R.ref_ts_set_timebins(ptr(timebin))
    act = np.sort(rng.choice(n, n // 2, replace=False)).astype(np.int32) if with_list else None
    na = n if act is None else len(act)
    for ti in (0, 1 << 5, 3 << 4, 1 << 11, (1 << 11) + 64):
        for mtb in (0, 3, 6, 46):
            out = np.zeros(na, np.int32)
            with quiet_stdout():
                m = R.ref_ts_active_sublist(ptr(act), c_i64(na), c_int(mtb), c_i64(ti), ptr(out))
            assert np.array_equal(o_active_sublist(act, na, timebin, skip, mtb, ti), out[:m]), (ti, mtb)
    vel = rng.standard_normal((n, 3)); acc = rng.standard_normal((n, 3)) * 100
    rvel = vel.copy()
    with quiet_stdout():
        R.ref_ts_hierarchical_kick(ptr(rvel), ptr(acc), ptr(act), c_i64(na), c_dbl(0.0371))
    assert np.array_equal(o_kick(vel, acc, act, skip, 0.0371), rvel)
