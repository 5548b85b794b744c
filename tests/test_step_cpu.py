"""SURVEY 8(f2) on the CPU: the oracle's drift / kick / active sub-list restatement against the reference's own
drift_all_particles (oracle/_ref, when built), the committed golden outputs of that code (tests/golden/step_ref.npz,
made by tests/golden/make_golden_step.py), and the literal semantics of libgadget/timestep.cpp:132-139, 962-968,
1353-1397.  No GPU."""
import os
import numpy as np
import pytest
from helpers import ROOT, oracle, ref, o_drift, o_kick, o_active_sublist, r_drift

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
