"""Test-side helpers: ctypes loaders for the checkers (oracle restatement, reference build) and the
canonical tree form used to compare topologies.  TEST INFRASTRUCTURE ONLY."""
import ctypes, os, subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libshenqi_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libshenqi_ref.so")

c_i64, c_dbl, c_int = ctypes.c_int64, ctypes.c_double, ctypes.c_int

# byte layout of the reference's struct NODE (libgadget/forcetree.h:38-66), 120 B
NODE_DTYPE = np.dtype([("sibling", "<i4"), ("father", "<i4"), ("len", "<f8"), ("center", "<f8", 3),
                       ("cofm", "<f8", 3), ("mass", "<f8"), ("hmax", "<f8"), ("suns", "<i4", 8),
                       ("noccupied", "<i4"), ("flags", "<u4")])
assert NODE_DTYPE.itemsize == 120


import contextlib


@contextlib.contextmanager
def quiet_stdout():
    """the reference's message() logs to stdout; silence fd 1 while it runs (bench.py must print one JSON line)."""
    import sys
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)
    try:
        yield
    finally:
        os.dup2(saved, 1)
        os.close(saved); os.close(devnull)


def ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


class GravParams(ctypes.Structure):
    """sqo_grav_params (oracle/shenqi_oracle.h)"""
    _fields_ = [("BoxSize", c_dbl), ("Nmesh", c_int), ("Asmth", c_dbl), ("G", c_dbl), ("ErrTolForceAcc", c_dbl),
                ("BHOpeningAngle", c_dbl), ("MaxBHOpeningAngle", c_dbl), ("TreeUseBH", c_int), ("TreeRcut", c_dbl),
                ("GravitySoftening", c_dbl), ("WindowType", c_int), ("rho0", c_dbl), ("update_potential", c_int)]


def default_params(box, nmesh, npart, **kw):
    p = dict(BoxSize=box, Nmesh=nmesh, Asmth=1.5, G=43.0071, ErrTolForceAcc=0.002, BHOpeningAngle=0.175,
             MaxBHOpeningAngle=0.9, TreeUseBH=0, TreeRcut=6.0, GravitySoftening=(box / round(npart ** (1 / 3.))) / 30.,
             WindowType=1, rho0=1.0, update_potential=0)
    p.update(kw)
    return p


def build_oracle():
    if not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "libshenqi_oracle.so"])


_oracle = None
def oracle():
    global _oracle
    if _oracle is None:
        build_oracle()
        L = ctypes.CDLL(ORACLE_SO)
        L.sqo_peano_hilbert_key.restype = ctypes.c_uint64
        L.sqo_tree_build.restype = c_i64
        L.sqo_tree_build_opt.restype = c_i64
        L.sqo_grav_short_tree.restype = c_i64
        L.sqo_drift.restype = c_i64
        L.sqo_is_timebin_active.argtypes = [c_int, c_i64]
        L.sqo_active_sublist.restype = c_i64
        L.sqo_gravity_timebins.restype = c_i64
        L.sqo_round_down_power_of_two.restype = c_i64
        L.sqo_round_down_power_of_two.argtypes = [c_i64]
        L.sqo_get_timestep_bin.argtypes = [c_i64]
        _oracle = L
    return _oracle


_ref = None
def ref():
    """The reference's own code (oracle/_ref); None when it was not built (no /root/reference at build time)."""
    global _ref
    if _ref is None:
        if not os.path.exists(REF_SO):
            return None
        L = ctypes.CDLL(REF_SO)
        L.ref_init()
        L.ref_peano_hilbert_key.restype = ctypes.c_uint64
        for f in ("ref_sort_particles", "ref_grav_short_tree", "ref_grav_short_pair", "ref_force_softening"):
            getattr(L, f).restype = c_dbl
        for f in ("ref_tree_build", "ref_tree_firstnode", "ref_tree_numnodes", "ref_grav_count_interactions", "ref_num_particles"):
            getattr(L, f).restype = c_i64
        _ref = L
    return _ref


# ----------------------------------------------------------------------------- oracle wrappers
def o_keys(pos, box):
    keys = np.zeros(len(pos), np.uint64)
    oracle().sqo_peano_keys(ptr(pos), c_i64(len(pos)), c_dbl(box), ptr(keys))
    return keys


def o_sort(keys, types=None):
    perm = np.zeros(len(keys), np.int64)
    oracle().sqo_sort_order(ptr(keys), ptr(types), c_i64(len(keys)), ptr(perm))
    return perm


def o_tree(pos, mass, box, firstnode=None, factor=1.5, shortcut=True):
    n = len(pos)
    firstnode = n if firstnode is None else firstnode
    maxnodes = int(factor * n) + 64
    nodes = np.zeros(maxnodes, NODE_DTYPE)
    nn = oracle().sqo_tree_build_opt(ptr(pos), ptr(mass), c_i64(n), c_dbl(box), ptr(nodes), c_i64(maxnodes), c_i64(firstnode),
                                     c_int(1 if shortcut else 0))
    assert nn > 0, "oracle tree build ran out of nodes"
    return nodes[:nn].copy(), firstnode


def o_walk(nodes, firstnode, pos, mass, oldacc, par, active=None, want_pot=False):
    n = len(pos)
    P = GravParams(**par)
    acc = np.zeros((n, 3)); pot = np.zeros(n) if want_pot else None; cnt = np.zeros(n, np.int64)
    act = None if active is None else np.ascontiguousarray(active, np.int32)
    tot = oracle().sqo_grav_short_tree(ptr(nodes), c_i64(firstnode), ptr(pos), ptr(mass), ptr(oldacc), ptr(act),
                                       c_i64(n if act is None else len(act)), ctypes.byref(P), ptr(acc), ptr(pot), ptr(cnt))
    return acc, pot, cnt, tot


def o_pair(pos, mass, par, want_pot=False):
    n = len(pos)
    P = GravParams(**par)
    acc = np.zeros((n, 3)); pot = np.zeros(n) if want_pot else None
    oracle().sqo_grav_short_pair(ptr(pos), ptr(mass), c_i64(n), ctypes.byref(P), ptr(acc), ptr(pot))
    return acc, pot


def o_cic_paint(pos, mass, box, nmesh):
    mesh = np.zeros((nmesh, nmesh, nmesh))
    oracle().sqo_cic_paint(ptr(pos), ptr(mass), c_i64(len(pos)), c_dbl(box), c_int(nmesh), ptr(mesh))
    return mesh


def o_cic_readout(mesh, pos, box, nmesh):
    out = np.zeros(len(pos))
    mesh = np.ascontiguousarray(mesh)
    oracle().sqo_cic_readout(ptr(mesh), ptr(pos), c_i64(len(pos)), c_dbl(box), c_int(nmesh), ptr(out))
    return out


def o_pm_force(pos, mass, box, nmesh, asmth, G):
    """Whole PM step restated: paint -> rfftn -> transfer -> irfftn (unnormalised, heFFTe scale::none,
    libgadget/petapm.cpp:54,66) -> readout.  Returns (potential, accel[N,3])."""
    mesh = o_cic_paint(pos, mass, box, nmesh)
    rho_k = np.ascontiguousarray(np.fft.rfftn(mesh))
    outs = [np.zeros_like(rho_k) for _ in range(4)]
    oracle().sqo_pm_transfer(ptr(rho_k), c_int(nmesh), c_dbl(box), c_dbl(asmth), c_dbl(G), *[ptr(o) for o in outs])
    res = []
    for o in outs:
        real = np.fft.irfftn(o, s=(nmesh,) * 3, axes=(0, 1, 2)) * float(nmesh) ** 3
        res.append(o_cic_readout(real, pos, box, nmesh))
    return res[0], np.stack(res[1:], 1)


def o_drift(pos, vel, skip, ddrift, shift, box):
    """sqo_drift on a copy; returns (new positions, number of non-finite coordinates)."""
    out = np.array(pos, np.float64, order="C", copy=True)
    vel = np.ascontiguousarray(vel, np.float64)
    skip = None if skip is None else np.ascontiguousarray(skip, np.uint8)
    shift = None if shift is None else np.ascontiguousarray(shift, np.float64)
    bad = oracle().sqo_drift(ptr(out), ptr(vel), ptr(skip), c_i64(len(out)), c_dbl(ddrift), ptr(shift), c_dbl(box))
    return out, int(bad)


def o_kick(vel, accel, active, skip, gravkick):
    out = np.array(vel, np.float64, order="C", copy=True)
    accel = np.ascontiguousarray(accel, np.float64)
    active = None if active is None else np.ascontiguousarray(active, np.int32)
    skip = None if skip is None else np.ascontiguousarray(skip, np.uint8)
    na = len(out) if active is None else len(active)
    oracle().sqo_kick(ptr(out), ptr(accel), ptr(active), ptr(skip), c_i64(na), c_dbl(gravkick))
    return out


def o_active_sublist(active, nactive, timebin, skip, maxtimebin, ti_current):
    active = None if active is None else np.ascontiguousarray(active, np.int32)
    timebin = np.ascontiguousarray(timebin, np.uint8)
    skip = None if skip is None else np.ascontiguousarray(skip, np.uint8)
    out = np.zeros(max(int(nactive), 1), np.int32)
    m = oracle().sqo_active_sublist(ptr(active), c_i64(nactive), ptr(timebin), ptr(skip), c_int(maxtimebin), c_i64(ti_current), ptr(out))
    return out[:m].copy()


class TimebinParams(ctypes.Structure):
    """sqo_timebin_params (same fields as the product's sq_timebin_params)."""
    _fields_ = [("atime", c_dbl), ("hubble", c_dbl), ("ErrTolIntAccuracy", c_dbl), ("MinSizeTimestep", c_dbl), ("softening", c_dbl),
                ("Ti_Current", c_i64), ("dti_max", c_i64), ("loga", c_dbl), ("lastsnap", c_i64), ("can_advance", c_int),
                ("sync_loga", c_dbl * 3), ("largest_active", c_int)]


def timebin_params(cls=TimebinParams, **kw):
    """A state like the reference's at z = 9 -> 8 with one sync point ahead: loga segments [ln 0.1, ln 0.111, ln 0.125]."""
    import math
    d = dict(atime=0.1, hubble=3.2, ErrTolIntAccuracy=0.02, MinSizeTimestep=0.0, softening=1.0 / 30, Ti_Current=0, dti_max=1 << 40,
             loga=math.log(0.1), lastsnap=0, can_advance=1, sync_loga=(math.log(0.1), math.log(0.111), math.log(0.125)), largest_active=40)
    d.update(kw)
    sl = d.pop("sync_loga")
    P = cls(**d)
    for k in range(3):
        P.sync_loga[k] = sl[k]
    return P


def o_gravity_timebins(P, accel, gravpm, active, skip, n, timebin=None):
    accel = np.ascontiguousarray(accel, np.float64); gravpm = np.ascontiguousarray(gravpm, np.float64)
    active = None if active is None else np.ascontiguousarray(active, np.int32)
    skip = None if skip is None else np.ascontiguousarray(skip, np.uint8)
    tb = np.full(n, 255, np.uint8) if timebin is None else np.array(timebin, np.uint8, copy=True)
    counts = np.zeros(47, np.int64)
    na = n if active is None else len(active)
    nbad = oracle().sqo_gravity_timebins(ctypes.byref(P), ptr(accel), ptr(gravpm), ptr(active), ptr(skip), c_i64(na), ptr(tb), ptr(counts))
    return tb, counts, int(nbad)


def o_direct27(pos, mass, box, G, h):
    """force_direct of libgadget/tests/test_gravity.cpp:133-153 (27 images, softened, self pairs skipped), in C."""
    pos = np.ascontiguousarray(pos, np.float64)
    mass = None if mass is None else np.ascontiguousarray(mass, np.float32)
    out = np.zeros((len(pos), 3))
    oracle().sqo_direct_sum_27(ptr(pos), ptr(mass), c_i64(len(pos)), c_dbl(box), c_dbl(G), c_dbl(h), ptr(out))
    return out


def boost_mt19937_uniform01(seed, count):
    """The numbers boost::random::uniform_real_distribution<double>(0, 1) draws from boost::random::mt19937(seed):
    boost 1.7x/1.8x generate_uniform_real() for an integer engine takes ONE 32-bit output per call,
    (x - min) / (max - min + 1) = x / 2^32 (boost/random/uniform_real_distribution.hpp, detail::generate_uniform_real,
    is_integral branch); mt19937(seed) is the standard init_genrand (first output for seed 0: 2357136044)."""
    bg = np.random.MT19937()
    bg._legacy_seeding(seed)
    raw = bg.random_raw(count)
    return raw.astype(np.float64) / 4294967296.0


def reference_gravity_test_samples():
    """The particle sets of libgadget/tests/test_gravity.cpp: (name, positions, check against the direct sum?).
    16^3 particles in a box of 8, Nmesh 48, Asmth 1.5, Rcut 7 (test_gravity.cpp:26-38, 197-247):
      flat    regular grid (:249-266), total force must vanish (:288-289)
      close   all particles within 16/5000 of (4,4,4) (:292-313)
      random0/1  two consecutive draws of do_random_test from ONE boost mt19937(0) stream (:316-355)."""
    box, n, nc = 8.0, 4096, 16
    i = np.arange(n)
    flat = np.stack([(box / nc) * (i // nc // nc), (box / nc) * ((i // nc) % nc), (box / nc) * (i % nc)], 1).astype(np.float64)
    close = np.stack([4. + (i // nc // nc) / 5000., 4. + ((i // nc) % nc) / 5000., 4. + (i % nc) / 5000.], 1).astype(np.float64)
    u = boost_mt19937_uniform01(0, 2 * 3 * n).reshape(2, n, 3)
    out = [("flat", flat, False), ("close", close, True)]
    a, b = n // 4, 3 * n // 4
    for k in range(2):
        pos = np.zeros((n, 3))
        pos[:a] = box * u[k, :a]
        pos[a:b] = box / 2 + box / 8 * np.exp((u[k, a:b] - 0.5) ** 2)
        pos[b:] = box * 0.1 + box / 32 * np.exp((u[k, b:] - 0.5) ** 2)
        out.append(("random%d" % k, pos, True))
    return out


def reference_gravity_test_metric(total, direct):
    """check_accns / find_means of libgadget/tests/test_gravity.cpp:85-131: component errors in units of the mean
    |component| of the computed force; returns (maxerr, meanerr)."""
    meanacc = np.abs(total).mean()
    err = np.abs(direct - total) / meanacc
    return float(err.max()), float(err.mean())


# ----------------------------------------------------------------------------- reference wrappers
def r_drift(pos, vel, flags, ddrift, shift, box):
    """The reference's drift_all_particles (libgadget/drift.cpp:81) on (pos, vel); flags bit0 IsGarbage, bit1 Swallowed."""
    R = ref()
    pos = np.ascontiguousarray(pos, np.float64)
    r_load(pos, np.ones(len(pos), np.float32), box)
    vel = np.ascontiguousarray(vel, np.float64)
    flags = None if flags is None else np.ascontiguousarray(flags, np.uint8)
    shift = np.ascontiguousarray(shift, np.float64)
    with quiet_stdout():
        R.ref_drift_all(ptr(vel), ptr(flags), c_dbl(ddrift), ptr(shift))
    out = np.zeros((len(pos), 3))
    R.ref_get_particles(ptr(out), None, None, None, None)
    return out


def r_load(pos, mass, box, ids=None):
    R = ref()
    ids = None if ids is None else np.ascontiguousarray(ids, np.uint64)
    R.ref_set_particles(ptr(pos), ptr(mass), ptr(ids), None, c_i64(len(pos)), c_dbl(box))


def r_sorted_particles():
    R = ref(); n = R.ref_num_particles()
    pos = np.zeros((n, 3)); mass = np.zeros(n, np.float32); ids = np.zeros(n, np.uint64)
    R.ref_get_particles(ptr(pos), ptr(mass), ptr(ids), None, None)
    return pos, mass, ids


def r_tree(factor=0.9):
    R = ref()
    secs = c_dbl()
    nn = R.ref_tree_build(c_dbl(factor), ctypes.byref(secs))
    nodes = np.zeros(nn, NODE_DTYPE)
    R.ref_tree_copy_nodes(ptr(nodes))
    return nodes, R.ref_tree_firstnode(), secs.value


def _r_args(par, meansep):
    return [c_int(par["Nmesh"]), c_dbl(par["Asmth"]), c_dbl(par["G"]), c_dbl(par["ErrTolForceAcc"]), c_dbl(par["BHOpeningAngle"]),
            c_dbl(par["MaxBHOpeningAngle"]), c_int(par["TreeUseBH"]), c_dbl(par["TreeRcut"]), c_dbl(par["GravitySoftening"] / meansep),
            c_dbl(meansep), c_int(par["WindowType"])]


def r_walk(par, meansep, fulltree=None, gravpm=None):
    R = ref(); n = R.ref_num_particles()
    if fulltree is not None or gravpm is not None:
        R.ref_set_accels(ptr(fulltree), ptr(gravpm))
    acc = np.zeros((n, 3))
    secs = R.ref_grav_short_tree(*_r_args(par, meansep), c_dbl(par["rho0"]), ptr(acc))
    return acc, secs


def r_counts(par, meansep):
    R = ref(); n = R.ref_num_particles()
    cnt = np.zeros(n, np.int64)
    tot = R.ref_grav_count_interactions(*_r_args(par, meansep), ptr(cnt))
    return cnt, tot


def r_pair(par, meansep):
    R = ref(); n = R.ref_num_particles()
    acc = np.zeros((n, 3))
    secs = R.ref_grav_short_pair(c_int(par["Nmesh"]), c_dbl(par["Asmth"]), c_dbl(par["G"]), c_dbl(par["TreeRcut"]),
                                 c_dbl(par["GravitySoftening"] / meansep), c_dbl(meansep), c_int(par["WindowType"]),
                                 c_dbl(par["rho0"]), ptr(acc))
    return acc, secs


# ----------------------------------------------------------------------------- the reference's own PM
def has_reference_pm():
    R = ref()
    return R is not None and hasattr(R, "ref_petapm_fft")


REF_SCRATCH = os.path.join(ROOT, "oracle", "_ref", "scratch")
UNIT_CM = 3.085678e21


def r_gravpm_force(pos, mass, box, nmesh, asmth, G, flags=None, types=None, dom=None, pot0=None):
    """gravpm_init_periodic + gravpm_force of the reference on (pos, mass); returns (Potential, GravPM, P(k) file rows)."""
    R = ref()
    n = len(pos)
    os.makedirs(REF_SCRATCH, exist_ok=True)
    out = os.path.join(REF_SCRATCH, "powerspectrum-0.1000.txt")
    if os.path.exists(out):
        os.remove(out)
    with quiet_stdout():
        r_load(pos, mass, box)
        if flags is not None:
            R.ref_set_flags(ptr(np.ascontiguousarray(flags, np.uint8)), ptr(None if types is None else np.ascontiguousarray(types, np.uint8)))
        R.ref_set_potential(ptr(np.zeros(n) if pot0 is None else np.ascontiguousarray(pot0, np.float64)))
        if dom is not None:
            R.ref_set_domain(ptr(dom.topnodes), c_int(len(dom.topnodes)), ptr(dom.topleaves), c_int(dom.ntopleaves))
        R.ref_set_power_dir(REF_SCRATCH.encode())
        try:
            R.ref_gravpm_force(c_int(nmesh), c_dbl(asmth), c_dbl(G))
        finally:
            if dom is not None:
                R.ref_set_domain(None, c_int(0), None, c_int(0))
    g = np.zeros((n, 3)); pot = np.zeros(n)
    R.ref_get_gravpm(ptr(g))
    R.ref_get_particles(None, None, None, None, ptr(pot))
    return pot, g, np.loadtxt(out, ndmin=2)


# ----------------------------------------------------------------------------- canonical tree form
def canonical_tree(nodes, firstnode, ids, box):
    """Node numbering and leaf-slot order are not canonical in the reference (thread caches, merge order),
    so compare the set of cells instead: for every node reachable from the root
       (level, ix, iy, iz, ChildType, TopLevel, InternalTopLevel, nparticles)   sorted,
    and for every leaf particle (level, ix, iy, iz, ID) sorted; plus per-cell (mass, cofm) in cell order."""
    N = nodes
    rootlen = N["len"][0]; corner = N["center"][0] - rootlen / 2
    frontier = np.array([0], np.int64)
    cells, parts, moms = [], [], []
    while len(frontier):
        nd = N[frontier]
        level = np.rint(np.log2(rootlen / nd["len"])).astype(np.int64)
        ijk = np.floor((nd["center"] - corner) / nd["len"][:, None]).astype(np.int64)
        ctype = (nd["flags"] >> 3) & 3
        top = (nd["flags"] >> 1) & 1; itop = nd["flags"] & 1
        isleaf = ctype == 0
        nocc = np.where(isleaf, nd["noccupied"], 0)
        cells.append(np.stack([level, ijk[:, 0], ijk[:, 1], ijk[:, 2], ctype, top, itop, nocc], 1))
        moms.append(np.concatenate([nd["mass"][:, None], nd["cofm"]], 1))
        lf = np.nonzero(isleaf)[0]
        for j in range(8):
            sel = lf[nd["noccupied"][lf] > j]
            if len(sel):
                pid = ids[nd["suns"][sel, j]].astype(np.int64)
                parts.append(np.stack([level[sel], ijk[sel, 0], ijk[sel, 1], ijk[sel, 2], pid], 1))
        inner = nd[ctype == 1]
        ch = inner["suns"].reshape(-1)
        frontier = ch[ch >= 0].astype(np.int64) - firstnode
    cells = np.concatenate(cells); moms = np.concatenate(moms)
    order = np.lexsort(cells[:, 3::-1].T)
    parts = np.concatenate(parts) if parts else np.zeros((0, 5), np.int64)
    porder = np.lexsort(parts[:, ::-1].T)
    return cells[order], parts[porder], moms[order]


# ----------------------------------------------------------------------------- synthetic particle sets
def random_clustered(n, box, seed=0):
    """3-component set of libgadget/tests/test_forcetree.cpp:404-430 (uniform + two clumps)."""
    rng = np.random.default_rng(seed)
    pos = np.zeros((n, 3))
    a, b = n // 4, 3 * n // 4
    pos[:a] = box * rng.random((a, 3))
    pos[a:b] = box / 2 + box / 8 * np.exp((rng.random((b - a, 3)) - 0.5) ** 2)
    pos[b:] = box * 0.1 + box / 32 * np.exp((rng.random((n - b, 3)) - 0.5) ** 2)
    return pos


def lattice(ncbrt, box):
    i = np.arange(ncbrt ** 3)
    return np.stack([(box / ncbrt) * (i // ncbrt // ncbrt), (box / ncbrt) * ((i // ncbrt) % ncbrt), (box / ncbrt) * (i % ncbrt)], 1).astype(np.float64)


def rel_err(a, b):
    return np.linalg.norm(a - b, axis=1) / np.maximum(np.linalg.norm(b, axis=1), 1e-300)


def direct_sum_27(pos, box, G, h):
    """periodic direct sum over 27 images with the spline softening of gravshort2.hpp:332-350
    (force_direct of libgadget/tests/test_gravity.cpp:133-153); unit masses."""
    n = len(pos)
    direct = np.zeros((n, 3))
    for sx in (-1, 0, 1):
        for sy in (-1, 0, 1):
            for sz in (-1, 0, 1):
                off = np.array([sx, sy, sz]) * box
                for i0 in range(0, n, 512):
                    d = off[None, None, :] + pos[i0:i0 + 512, None, :] - pos[None, :, :]
                    r2 = (d ** 2).sum(-1); r = np.sqrt(r2)
                    with np.errstate(divide="ignore", invalid="ignore"):
                        fac = 1 / (r2 * r)
                        u = r / h
                        f1 = (10.666666666667 + u * u * (32.0 * u - 38.4)) / h ** 3
                        f2 = (21.333333333333 - 48.0 * u + 38.4 * u * u - 10.666666666667 * u ** 3 - 0.066666666667 / u ** 3) / h ** 3
                    fac = np.where(r < h, np.where(u < 0.5, f1, f2), fac)
                    fac[r2 == 0] = 0
                    direct[i0:i0 + 512] += -(d * fac[..., None]).sum(1) * G
    return direct
