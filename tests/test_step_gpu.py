"""SURVEY 8(f2) on the GPU, through the C ABI: sq_drift_device / sq_kick_device / sq_active_sublist_device against the
oracle restatement (bit for bit: FP64 with separately rounded products and sums, integer lists), the committed golden
outputs of the reference's own drift_all_particles (tests/golden/step_ref.npz), and size-independent properties at the
full 256^3 load of BASELINE config 2."""
import os
import sys
import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE); sys.path.insert(0, os.path.dirname(HERE))
from helpers import o_drift, o_kick, o_active_sublist, o_gravity_timebins, timebin_params

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(HERE, "golden", "step_ref.npz"))


@pytest.fixture(scope="module")
def ctx():
    import shenqi_b200 as sq
    c = sq.Context(0)
    c.set_stream(torch.cuda.current_stream().cuda_stream)
    yield c
    torch.cuda.synchronize()
    c.close()


def dev(a):
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).cuda()


def dptr(t):
    return None if t is None else t.data_ptr()


def gpu_drift(ctx, pos, vel, skip, ddrift, shift, box):
    p, v, s = dev(pos), dev(vel), dev(skip)
    ctx.drift_device(p.data_ptr(), v.data_ptr(), len(pos), ddrift, shift, box, skip_ptr=dptr(s))
    return p.cpu().numpy()


def test_drift_equals_golden_reference_output(ctx):
    pos, vel, flags, ddrift, shift, box = GOLD["pos"], GOLD["vel"], GOLD["flags"], float(GOLD["ddrift"]), GOLD["shift"], float(GOLD["box"])
    skip = (flags != 0).astype(np.uint8)
    assert np.array_equal(gpu_drift(ctx, pos, vel, skip, ddrift, shift, box), GOLD["pos_after"])
    assert np.array_equal(gpu_drift(ctx, pos, vel, None, -ddrift, None, box), GOLD["pos_after_back"])


@pytest.mark.parametrize("n", [1, 33, 100003, 1 << 20])
def test_drift_equals_oracle(ctx, n):
    rng = np.random.default_rng(n)
    box = 250000.0
    pos = rng.random((n, 3)) * box
    pos.flat[::7] = box                                   # Pos == BoxSize stays (wrap is `> Box`, drift.cpp:77)
    vel = rng.standard_normal((n, 3)) * 2e5
    vel[rng.random(n) < 0.01] *= 30
    skip = (rng.random(n) < 0.04).astype(np.uint8)
    shift = np.array([1234.5, -99999.25, 0.0])
    exp, bad = o_drift(pos, vel, skip, 0.83, shift, box)
    assert bad == 0
    got = gpu_drift(ctx, pos, vel, skip, 0.83, shift, box)
    assert np.array_equal(got, exp)
    assert got.min() > 0 and got.max() <= box


def test_drift_empty_and_nonfinite(ctx):
    import shenqi_b200 as sq
    ctx.drift_device(0, 0, 0, 1.0, None, 10.0)            # n = 0: nothing to do
    pos = np.full((5, 3), 1.0); vel = np.zeros((5, 3)); vel[3, 2] = np.inf
    with pytest.raises(sq.ShenqiError, match="non-finite"):
        gpu_drift(ctx, pos, vel, None, 1.0, None, 10.0)    # endrun(5) in the reference, drift.cpp:70-73
    with pytest.raises(sq.ShenqiError):
        gpu_drift(ctx, pos, np.zeros((5, 3)), None, 1.0, None, 0.0)


@pytest.mark.parametrize("with_list", [False, True])
def test_kick_equals_oracle(ctx, with_list):
    rng = np.random.default_rng(11)
    n = 300007
    vel = rng.standard_normal((n, 3)) * 100; acc = rng.standard_normal((n, 3)) * 1e4
    skip = (rng.random(n) < 0.05).astype(np.uint8)
    act = np.sort(rng.choice(n, n // 5, replace=False)).astype(np.int32) if with_list else None
    na = n if act is None else len(act)
    v, a, s, l = dev(vel), dev(acc), dev(skip), dev(act)
    ctx.kick_device(v.data_ptr(), a.data_ptr(), na, 0.00271, active_ptr=dptr(l), skip_ptr=s.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(v.cpu().numpy(), o_kick(vel, acc, act, skip, 0.00271))
    ctx.kick_device(v.data_ptr(), a.data_ptr(), 0, 1.0)   # empty active set


@pytest.mark.parametrize("n", [1, 31, 2047, 2048, 2049, 10241, 3000017])
@pytest.mark.parametrize("with_list", [False, True])
def test_active_sublist_equals_oracle(ctx, n, with_list):
    rng = np.random.default_rng(n + with_list)
    timebin = rng.integers(0, 10, n).astype(np.uint8)
    skip = (rng.random(n) < 0.03).astype(np.uint8)
    act = np.sort(rng.choice(n, max(1, n // 2), replace=False)).astype(np.int32) if with_list else None
    na = n if act is None else len(act)
    tb, s, l = dev(timebin), dev(skip), dev(act)
    out = torch.empty(na, dtype=torch.int32, device="cuda")
    for ti, mtb in ((0, 46), (1 << 4, 46), (1 << 4, 2), (3 << 3, 5), ((1 << 9) + 1, 46), (1 << 9, 0)):
        m = ctx.build_active_sublist_device(na, tb.data_ptr(), mtb, ti, out.data_ptr(), active_ptr=dptr(l), skip_ptr=s.data_ptr())
        exp = o_active_sublist(act, na, timebin, skip, mtb, ti)
        assert m == len(exp)
        assert np.array_equal(out[:m].cpu().numpy(), exp)
    assert ctx.build_active_sublist_device(0, tb.data_ptr(), 46, 0, out.data_ptr()) == 0


def test_full_size_properties(ctx):
    """256^3 particles (BASELINE config 2): the drift against the same IEEE operations done by torch elementwise kernels,
    positions inside (0, Box], zero drift = identity; the sub-list is strictly increasing, its length equals the
    predicate count, selecting from it again changes nothing (idempotence), and nested bins give nested lists."""
    n, box = 256 ** 3, 100000.0
    g = torch.Generator(device="cuda").manual_seed(3)
    pos = torch.rand((n, 3), dtype=torch.float64, device="cuda", generator=g) * box
    pos.clamp_(min=1e-9)
    vel = torch.randn((n, 3), dtype=torch.float64, device="cuda", generator=g) * 4e4
    shift = torch.tensor([17.0, -3.25, 50000.0], dtype=torch.float64, device="cuda")
    exp = (pos + shift) + vel * 1.7
    for _ in range(8):
        exp = torch.where(exp > box, exp - box, exp)
    for _ in range(8):
        exp = torch.where(exp <= 0, exp + box, exp)
    p = pos.clone()
    ctx.drift_device(p.data_ptr(), vel.data_ptr(), n, 1.7, shift.tolist(), box)
    assert torch.equal(p, exp)
    assert float(p.min()) > 0 and float(p.max()) <= box
    q = p.clone()
    ctx.drift_device(q.data_ptr(), vel.data_ptr(), n, 0.0, None, box)
    assert torch.equal(q, p)
    del exp, q, pos

    acc = torch.randn((n, 3), dtype=torch.float64, device="cuda", generator=g)
    v2 = vel.clone()
    ctx.kick_device(v2.data_ptr(), acc.data_ptr(), n, 0.125)
    torch.cuda.synchronize()
    assert torch.equal(v2, vel + acc * 0.125)
    del v2, acc

    timebin = torch.randint(0, 12, (n,), dtype=torch.uint8, device="cuda", generator=g)
    out = torch.empty(n, dtype=torch.int32, device="cuda")
    ti = 3 << 6
    m = ctx.build_active_sublist_device(n, timebin.data_ptr(), 8, ti, out.data_ptr())
    tb = timebin.long()
    pred = (tb <= 8) & ((tb == 0) | (torch.remainder(torch.full_like(tb, ti), torch.bitwise_left_shift(torch.ones_like(tb), tb)) == 0))
    assert m == int(pred.sum())
    sub = out[:m]
    assert bool((sub[1:] > sub[:-1]).all())
    assert torch.equal(sub.long(), torch.nonzero(pred).flatten())
    out2 = torch.empty(m, dtype=torch.int32, device="cuda")
    m2 = ctx.build_active_sublist_device(m, timebin.data_ptr(), 8, ti, out2.data_ptr(), active_ptr=sub.data_ptr())
    assert m2 == m and torch.equal(out2, sub)
    m3 = ctx.build_active_sublist_device(m, timebin.data_ptr(), 4, ti, out2.data_ptr(), active_ptr=sub.data_ptr())
    assert 0 < m3 < m and bool((timebin[out2[:m3].long()] <= 4).all())


@pytest.mark.parametrize("with_list", [False, True])
def test_gravity_timebins_equal_oracle(ctx, with_list):
    """timestep.cpp:351-377: bins and timebincounts identical to the restatement (integer result of an FP64 chain whose
    every operation is IEEE-rounded in source order on both sides)."""
    import shenqi_b200 as sq
    rng = np.random.default_rng(21)
    n = 500009
    acc = rng.standard_normal((n, 3)) * 10.0 ** rng.uniform(-2, 6, (n, 1))
    pm = rng.standard_normal((n, 3)) * 10.0
    acc[:3] = 0; pm[:3] = 0
    skip = (rng.random(n) < 0.05).astype(np.uint8)
    act = np.sort(rng.choice(n, n // 3, replace=False)).astype(np.int32) if with_list else None
    na = n if act is None else len(act)
    a, g, s, l = dev(acc), dev(pm), dev(skip), dev(act)
    ti = (1 << 45) + (1 << 30)
    for kw in (dict(), dict(Ti_Current=ti, loga=np.log(0.1) + ti * (np.log(0.111) - np.log(0.1)) / (1 << 46)), dict(largest_active=33),
               dict(dti_max=1 << 31), dict(MinSizeTimestep=1e-4), dict(can_advance=0), dict(dti_max=0)):
        exp_tb, exp_counts, exp_bad = o_gravity_timebins(timebin_params(**kw), acc, pm, act, skip, n)
        tb = torch.full((n,), 255, dtype=torch.uint8, device="cuda")
        counts, nbad = ctx.gravity_timebins_device(timebin_params(cls=sq.TimebinParams, **kw), a.data_ptr(), g.data_ptr(), na, tb.data_ptr(),
                                                   active_ptr=dptr(l), skip_ptr=s.data_ptr())
        assert np.array_equal(tb.cpu().numpy(), exp_tb)
        assert counts == exp_counts.tolist() and nbad == exp_bad
    # push-down (timestep.cpp:399-407) with the host rule on the counts
    P = timebin_params(cls=sq.TimebinParams)
    tb = torch.full((n,), 255, dtype=torch.uint8, device="cuda")
    counts, _ = ctx.gravity_timebins_device(P, a.data_ptr(), g.data_ptr(), na, tb.data_ptr(), active_ptr=dptr(l), skip_ptr=s.data_ptr())
    before = tb.cpu().numpy()
    occupied = [b for b in range(47) if counts[b]]
    push = occupied[len(occupied) // 2]
    ctx.timebin_pushdown_device(na, tb.data_ptr(), push, active_ptr=dptr(l))
    torch.cuda.synchronize()
    exp = before.copy()
    idx = np.arange(n) if act is None else act
    exp[idx] = np.minimum(before[idx], push)                 # the reference clamps every listed particle, skipped or not
    assert np.array_equal(tb.cpu().numpy(), exp)
    pd, merged = sq.pushdown_bin(counts, 40, True)
    assert 0 < pd <= max(occupied) and sum(merged[:pd + 1]) >= sum(counts) - sum(counts[pd + 1:])
    ctx.gravity_timebins_device(P, 0, 0, 0, 0)               # empty active set


def test_hierarchical_cycle_at_full_size(ctx):
    """256^3: assign bins -> counts sum to N -> sub-lists per bin partition the particles the way the counts say -> kick of a
    sub-list touches exactly its members (a whole hierarchical bookkeeping cycle, timestep.cpp:351-472, on the device)."""
    import shenqi_b200 as sq
    n = 256 ** 3
    g = torch.Generator(device="cuda").manual_seed(5)
    acc = torch.randn((n, 3), dtype=torch.float64, device="cuda", generator=g) * torch.pow(10.0, torch.rand((n, 1), dtype=torch.float64, device="cuda", generator=g) * 6 - 1)
    pm = torch.randn((n, 3), dtype=torch.float64, device="cuda", generator=g)
    tb = torch.empty(n, dtype=torch.uint8, device="cuda")
    P = timebin_params(cls=sq.TimebinParams)
    counts, nbad = ctx.gravity_timebins_device(P, acc.data_ptr(), pm.data_ptr(), n, tb.data_ptr())
    assert nbad == 0 and sum(counts) == n
    assert counts == torch.bincount(tb.long(), minlength=47).tolist()
    occupied = [b for b in range(47) if counts[b]]
    assert len(occupied) >= 4
    out = torch.empty(n, dtype=torch.int32, device="cuda")
    vel = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
    b = occupied[1]
    m = ctx.build_active_sublist_device(n, tb.data_ptr(), b, 1 << b, out.data_ptr())
    assert m == sum(counts[:b + 1])                          # Ti = 2^b: exactly the bins <= b are active
    ctx.kick_device(vel.data_ptr(), acc.data_ptr(), m, 0.5, active_ptr=out.data_ptr())
    torch.cuda.synchronize()
    touched = (vel != 0).any(dim=1)
    assert int(touched.sum()) == m and bool((tb[touched] <= b).all())
    assert torch.equal(vel[out[:m].long()], acc[out[:m].long()] * 0.5)


def test_substep_with_device_resident_list():
    """One hierarchical sub-step without the list ever visiting the host (timestep.cpp:425-472): bins -> sub-list (device)
    -> tree of the sub-list -> walk of the sub-list -> kick through the sub-list.  Same tree, interaction count and
    accelerations (bit for bit) as the same sub-step driven by a host copy of the list."""
    import shenqi_b200 as sq
    from helpers import random_clustered
    box, n, nmesh = 8.0, 60000, 48
    pos = random_clustered(n, box, 17); mass = np.ones(n, np.float32)
    P = sq.Params.defaults(box, nmesh, (box / n ** (1 / 3.)) / 30., TreeUseBH=1, TreeRcut=6.0)
    c = sq.Context(0)
    try:
        c.set_params(P)
        c.upload_particles(pos, mass); c.ph_sort()
        rng = np.random.default_rng(3)
        timebin = dev(rng.integers(1, 6, n).astype(np.uint8))
        out = torch.empty(n, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        m = c.build_active_sublist_device(n, timebin.data_ptr(), 3, 1 << 2, out.data_ptr())     # bins 1 and 2 are active at Ti = 4
        c._check(c.L.sq_sync(c._ctx))
        host_list = out[:m].cpu().numpy()
        assert 0 < m < n and np.array_equal(host_list, np.nonzero(timebin.cpu().numpy() <= 2)[0])
        info_h = c.tree_build(host_list)
        st_h = c.grav_short_tree_device(host_list)
        acc_h, _ = c.download_accel()
        info_d = c.tree_build_device_list(out.data_ptr(), m)
        st_d = c.grav_short_tree_device_list(out.data_ptr(), m)
        acc_d, _ = c.download_accel()
        assert info_d.numparticles == info_h.numparticles == m and info_d.numnodes == info_h.numnodes
        assert st_d.interactions == st_h.interactions > 0
        assert np.array_equal(acc_d, acc_h) and np.abs(acc_d[host_list]).max() > 0
        a = dev(acc_d); vel = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        c.kick_device(vel.data_ptr(), a.data_ptr(), m, 0.25, active_ptr=out.data_ptr())
        c._check(c.L.sq_sync(c._ctx))
        exp = np.zeros((n, 3)); exp[host_list] = acc_d[host_list] * 0.25
        assert np.array_equal(vel.cpu().numpy(), exp)
    finally:
        c.close()
