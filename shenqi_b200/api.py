"""ctypes binding of include/shenqi_b200.h, shaped after the reference's entry points
(slots_gc_sorted, force_tree_full / force_tree_active_moments, grav_short_tree, gravpm_force)."""
import ctypes
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
c_i64, c_dbl, c_int = ctypes.c_int64, ctypes.c_double, ctypes.c_int

# struct NODE, libgadget/forcetree.h:38-66 (120 B)
NODE_DTYPE = np.dtype([("sibling", "<i4"), ("father", "<i4"), ("len", "<f8"), ("center", "<f8", 3),
                       ("cofm", "<f8", 3), ("mass", "<f8"), ("hmax", "<f8"), ("suns", "<i4", 8),
                       ("noccupied", "<i4"), ("flags", "<u4")])
# struct particle_data, libgadget/partmanager.h:9-71 (160 B, LOW_PRECISION=double, non-DEBUG)
PARTICLE_DTYPE = np.dtype({"names": ["Pos", "TopLeaf", "Mass", "PI", "flags", "TimeBinHydro", "TimeBinGravity", "Type",
                                     "Vel", "FullTreeGravAccel", "GravPM", "Ti_drift", "Hsml", "DtHsml", "ID", "GrNr", "Potential"],
                           "formats": [("<f8", 3), "<i4", "<f4", "<i4", "u1", "u1", "u1", "u1",
                                       ("<f8", 3), ("<f8", 3), ("<f8", 3), "<i8", "<f8", "<f8", "<u8", "<i8", "<f8"],
                           "offsets": [0, 24, 28, 32, 36, 37, 38, 39, 40, 64, 88, 112, 120, 128, 136, 144, 152],
                           "itemsize": 160})


class ShenqiError(RuntimeError):
    pass


class Params(ctypes.Structure):
    """sq_params: the parameter-file knobs that reach the path (gadget/params.cpp:29-33, 84-104, 161-165)."""
    _fields_ = [("BoxSize", c_dbl), ("Nmesh", c_int), ("Asmth", c_dbl), ("G", c_dbl), ("ErrTolForceAcc", c_dbl),
                ("BHOpeningAngle", c_dbl), ("MaxBHOpeningAngle", c_dbl), ("TreeUseBH", c_int), ("TreeRcut", c_dbl),
                ("GravitySoftening", c_dbl), ("WindowType", c_int), ("rho0", c_dbl), ("TreeAllocFactor", c_dbl)]

    @classmethod
    def defaults(cls, BoxSize, Nmesh, GravitySoftening, **kw):
        p = cls(BoxSize=BoxSize, Nmesh=Nmesh, Asmth=1.5, G=43.0071, ErrTolForceAcc=0.002, BHOpeningAngle=0.175,
                MaxBHOpeningAngle=0.9, TreeUseBH=2, TreeRcut=6.0, GravitySoftening=GravitySoftening, WindowType=1,
                rho0=1.0, TreeAllocFactor=0.9)
        for k, v in kw.items():
            if not hasattr(p, k):
                raise KeyError(k)
            setattr(p, k, v)
        return p


class AosLayout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_size_t) for n in ("stride", "off_pos", "off_mass", "off_flags", "off_fulltreegravaccel",
                                               "off_gravpm", "off_id", "off_potential")]


class WalkStats(ctypes.Structure):
    _fields_ = [("ntargets", c_i64), ("ngroups", c_i64), ("interactions", c_i64), ("list_entries", c_i64),
                ("nodes_tested", c_i64), ("pair_tests", c_i64), ("exact_tests", c_i64), ("kernel_ms", c_dbl), ("pseudo_opened", c_i64)]


class TreeInfo(ctypes.Structure):
    _fields_ = [("numnodes", c_i64), ("numparticles", c_i64), ("maxnodes", c_i64), ("maxlevel", c_int), ("geo_mismatch", c_int)]


class LetInfo(ctypes.Structure):
    """sq_let_info: what sq_tree_build_dist built, imported and exported."""
    _fields_ = [(n, c_i64) for n in ("local_nodes", "local_particles", "imported_nodes", "imported_particles", "exported_nodes",
                                     "exported_particles", "bytes_sent", "bytes_received", "toptree_exports")] + \
               [(n, c_dbl) for n in ("build_ms", "moments_exchange_ms", "let_mark_ms", "let_exchange_ms", "let_attach_ms")]


# struct topnode_data / topleaf_data / task_data, libgadget/domain.h:12-31 (24 / 12 / 8 bytes)
TOPNODE_DTYPE = np.dtype({"names": ["StartKey", "Daughter", "Shift", "Leaf"], "formats": ["<u8", "<i4", "<i4", "<i4"], "offsets": [0, 8, 12, 16], "itemsize": 24})
TOPLEAF_DTYPE = np.dtype([("Task", "<i4"), ("topnode", "<i4"), ("treenode", "<i4")])
TASK_DTYPE = np.dtype([("StartLeaf", "<i4"), ("EndLeaf", "<i4")])
NCCL_UNIQUE_ID_BYTES = 128


class Domain:
    """The reference's DomainDecomp (libgadget/domain.h:33-43) as three arrays."""

    def __init__(self, topnodes, topleaves, tasks):
        self.topnodes = np.ascontiguousarray(topnodes, TOPNODE_DTYPE)
        self.topleaves = np.ascontiguousarray(topleaves, TOPLEAF_DTYPE)
        self.tasks = np.ascontiguousarray(tasks, TASK_DTYPE)

    @property
    def ntopleaves(self):
        return len(self.topleaves)

    def validate(self):
        """sq_domain_validate (pure host code): raises if sq_domain_set would refuse this domain; returns the depth of
        the deepest top node."""
        L = load_library()
        ml = c_int()
        rc = L.sq_domain_validate(_p(self.topnodes), c_int(len(self.topnodes)), _p(self.topleaves), c_int(len(self.topleaves)),
                                  _p(self.tasks), c_int(len(self.tasks)), ctypes.byref(ml))
        if rc != 0:
            raise ShenqiError("shenqi_b200 error %d: %s" % (rc, L.sq_last_error().decode()))
        return ml.value

    def topleaf_of(self, keys):
        """domain_get_topleaf (domain.h:68-75), vectorised."""
        keys = np.asarray(keys, np.uint64)
        no = np.zeros(len(keys), np.int64)
        T = self.topnodes
        while True:
            d = T["Daughter"][no]
            live = d >= 0
            if not live.any():
                break
            sub = ((keys[live] - T["StartKey"][no[live]]) >> (T["Shift"][no[live]] - 3).astype(np.uint64)).astype(np.int64)
            no[live] = d[live] + sub
        return T["Leaf"][no]


def domain_assign_balanced(cost, nranks, nsegment_per_task=1):
    """sq_domain_assign_balanced (pure host code): the reference's own assignment of top leaves to tasks
    (domain_assign_topleaves_balanced, libgadget/domain.cpp:620-778); cost = integer cost per top leaf in key order;
    returns the task of every leaf."""
    L = load_library()
    cost = np.ascontiguousarray(cost, np.int64)
    task = np.zeros(len(cost), np.int32)
    rc = L.sq_domain_assign_balanced(_p(cost), c_int(len(cost)), c_int(nranks), c_int(nsegment_per_task), _p(task))
    if rc != 0:
        raise ShenqiError("shenqi_b200 error %d: %s" % (rc, L.sq_last_error().decode()))
    return task


def domain_from_histogram(count, nranks, overdecomposition=4, cost=None, max_topnodes=None, assign="bisect"):
    """sq_domain_from_histogram (pure host code, no GPU needed): a Domain from a Peano-Hilbert key histogram over the
    8^level cells of one level.  assign = "bisect": segments cut where the running cost crosses k / nranks of the total;
    "reference": the leaves of the same top tree assigned by the reference's own greedy rule (domain_assign_balanced;
    costs rounded to integers, as the reference's are)."""
    D = _domain_from_histogram(count, nranks, overdecomposition, cost, max_topnodes)
    if assign == "bisect":
        return D
    assert assign == "reference", assign
    count = np.ascontiguousarray(count, np.int64)
    w = count if cost is None else np.rint(np.ascontiguousarray(cost, np.float64)).astype(np.int64)
    level = int(round(np.log2(len(count)) / 3))
    T = D.topnodes[D.topleaves["topnode"]]                   # leaves are in key order
    a = (T["StartKey"] >> np.uint64(63 - 3 * level)).astype(np.int64)
    b = a + (np.int64(1) << (T["Shift"].astype(np.int64) - (63 - 3 * level)))
    pre = np.concatenate([[0], np.cumsum(w)])
    task = domain_assign_balanced(pre[b] - pre[a], nranks)
    D.topleaves["Task"] = task
    for t in range(nranks):
        mine = np.flatnonzero(task == t)
        D.tasks[t] = (mine[0], mine[-1] + 1)
    return D


def _domain_from_histogram(count, nranks, overdecomposition=4, cost=None, max_topnodes=None):
    L = load_library()
    count = np.ascontiguousarray(count, np.int64)
    level = int(round(np.log2(len(count)) / 3))
    assert 8 ** level == len(count)
    cost = None if cost is None else np.ascontiguousarray(cost, np.float64)
    mx = max_topnodes or (8 * len(count) // 7 + 16)
    tn = np.zeros(mx, TOPNODE_DTYPE); tl = np.zeros(mx, TOPLEAF_DTYPE); tk = np.zeros(nranks, TASK_DTYPE)
    ntn, ntl = c_int(), c_int()
    rc = L.sq_domain_from_histogram(_p(count), _p(cost), c_int(level), c_int(nranks), c_int(overdecomposition), _p(tn), _p(tl), _p(tk), c_int(mx),
                                    ctypes.byref(ntn), ctypes.byref(ntl))
    if rc != 0:
        raise ShenqiError("shenqi_b200 error %d: %s" % (rc, L.sq_last_error().decode()))
    return Domain(tn[:ntn.value].copy(), tl[:ntl.value].copy(), tk)


def bigfile_block_info(path, block):
    """(dtype, nmemb, nrows, nfile) of a bigfile block (sq_bigfile_block_info; host code, no GPU needed)."""
    L = load_library()
    dt = ctypes.create_string_buffer(8); nm, nf, nr = c_int(), c_int(), c_i64()
    rc = L.sq_bigfile_block_info(path.encode(), block.encode(), dt, ctypes.byref(nm), ctypes.byref(nr), ctypes.byref(nf))
    if rc != 0:
        raise ShenqiError("shenqi_b200 error %d: %s" % (rc, L.sq_last_error().decode()))
    return dt.value.decode(), nm.value, nr.value, nf.value


def bigfile_read_block(path, block, first=0, count=None, want="f8"):
    """rows [first, first + count) of a bigfile block as a numpy array [count, nmemb] of `want`."""
    L = load_library()
    _, nmemb, nrows, _ = bigfile_block_info(path, block)
    count = nrows - first if count is None else count
    out = np.zeros((count, nmemb), {"f8": np.float64, "f4": np.float32, "u8": np.uint64, "i8": np.int64, "u4": np.uint32, "i4": np.int32, "u1": np.uint8}[want])
    rc = L.sq_bigfile_read_block(path.encode(), block.encode(), c_i64(first), c_i64(count), _p(out), want.encode())
    if rc != 0:
        raise ShenqiError("shenqi_b200 error %d: %s" % (rc, L.sq_last_error().decode()))
    return out


def bigfile_read_attr(path, block, name, maxn=64):
    L = load_library()
    v = np.zeros(maxn); n = c_int()
    rc = L.sq_bigfile_read_attr(path.encode(), block.encode(), name.encode(), _p(v), c_int(maxn), ctypes.byref(n))
    if rc != 0:
        raise ShenqiError("shenqi_b200 error %d: %s" % (rc, L.sq_last_error().decode()))
    return v[:min(n.value, maxn)].copy()


def comm_unique_id():
    """ncclGetUniqueId as 128 bytes (numpy uint8): create on one rank, broadcast to all."""
    L = load_library()
    buf = np.zeros(NCCL_UNIQUE_ID_BYTES, np.uint8)
    rc = L.sq_comm_unique_id(_p(buf))
    if rc != 0:
        raise ShenqiError("shenqi_b200 error %d: %s" % (rc, L.sq_last_error().decode()))
    return buf


def comm_loopback_id(key=0):
    """TEST TRANSPORT id (sq_comm_loopback_id): ranks as threads of this process meet in an in-process hub."""
    L = load_library()
    buf = np.zeros(NCCL_UNIQUE_ID_BYTES, np.uint8)
    L.sq_comm_loopback_id(_p(buf), c_int(key))
    return buf


class Timings(ctypes.Structure):
    _fields_ = [(n, c_dbl) for n in ("keys_ms", "sort_ms", "gather_ms", "tree_build_ms", "tree_moments_ms", "walk_ms",
                                     "pm_paint_ms", "pm_fft_ms", "pm_transfer_ms", "pm_readout_ms")] + [("kernel_launches", c_i64)]


class TimebinParams(ctypes.Structure):
    """sq_timebin_params: the scalars of the gravity time-bin assignment (libgadget/timestep.cpp:351-377)."""
    _fields_ = [("atime", c_dbl), ("hubble", c_dbl), ("ErrTolIntAccuracy", c_dbl), ("MinSizeTimestep", c_dbl), ("softening", c_dbl),
                ("Ti_Current", c_i64), ("dti_max", c_i64), ("loga", c_dbl), ("lastsnap", c_i64), ("can_advance", c_int),
                ("sync_loga", c_dbl * 3), ("largest_active", c_int)]


TIMEBINS = 46

EXPORTS = ["sq_last_error", "sq_version", "sq_ctx_create", "sq_ctx_destroy", "sq_set_params", "sq_set_walk_raw", "sq_sync", "sq_default_aos_layout",
           "sq_upload_particles_aos", "sq_upload_particles_soa", "sq_set_old_accel", "sq_set_oldacc", "sq_ph_sort", "sq_get_keys",
           "sq_get_sorted", "sq_tree_build", "sq_tree_export_nodes", "sq_tree_walk", "sq_tree_walk_device", "sq_download_accel",
           "sq_pm_force", "sq_pm_force_device", "sq_gravity_step", "sq_pm_paint", "sq_pm_get_mesh", "sq_get_timings", "sq_reset_timings", "sq_measure_fp32_peak",
           "sq_ctx_set_stream", "sq_device_ptr", "sq_upload_particles_device", "sq_get_geo_mismatch", "sq_set_sorted_device", "sq_set_oldacc_range_device", "sq_tree_walk_range",
           "sq_pm_slab_route", "sq_pm_slab_unroute", "sq_pm_slab_paint", "sq_pm_slab_fft_fwd", "sq_pm_slab_fft_x", "sq_pm_slab_transfer", "sq_pm_slab_transfer_phi", "sq_pm_slab_gradient",
           "sq_pm_slab_fft_inv", "sq_pm_slab_readout", "sq_drift_device", "sq_kick_device", "sq_active_sublist_device",
           "sq_gravity_timebins_device", "sq_timebin_pushdown_device",
           "sq_comm_unique_id", "sq_comm_loopback_id", "sq_ctx_create_dist", "sq_comm_rank", "sq_domain_set", "sq_domain_get_treenodes", "sq_domain_from_histogram", "sq_domain_assign_balanced", "sq_domain_validate",
           "sq_domain_histogram", "sq_domain_owner", "sq_tree_build_dist", "sq_toptree_export_count", "sq_download_accel_device",
           "sq_walk_particle_cost", "sq_keys_device", "sq_tree_export_nodes_ex", "sq_tree_local_nodes",
           "sq_bigfile_block_info", "sq_bigfile_read_block", "sq_bigfile_read_attr", "sq_upload_particles_bigfile"]

PTR_SORTED_KEYS, PTR_SORTED_POS, PTR_SORTED_MASS, PTR_SORTED_PERM, PTR_WALK_ACCEL, PTR_WALK_POT, PTR_PM_OUT = range(7)


def library_path():
    return os.path.join(_HERE, "libshenqi_b200.so")


_lib = None
def load_library():
    """Load libshenqi_b200.so.  Raises (loudly) if it has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            raise ShenqiError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(or make -C shenqi_b200/csrc). shenqi_b200 has no CPU fallback." % path)
        L = ctypes.CDLL(path)
        L.sq_last_error.restype = ctypes.c_char_p
        L.sq_version.restype = ctypes.c_char_p
        L.sq_ctx_destroy.restype = None
        L.sq_default_aos_layout.restype = None
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


class Context:
    """One GPU context (sq_ctx).  Method names follow the reference functions they stand behind."""

    def __init__(self, device=0, nccl_id=None, rank=0, nranks=1):
        """nranks > 1: one rank of a multi-GPU run (sq_ctx_create_dist, collective); nccl_id = comm_unique_id() of one rank."""
        self.L = load_library()
        self._ctx = ctypes.c_void_p()
        self.rank, self.nranks = rank, nranks
        if nranks > 1 or nccl_id is not None:
            idb = None if nccl_id is None else np.ascontiguousarray(nccl_id, np.uint8)
            self._check(self.L.sq_ctx_create_dist(ctypes.byref(self._ctx), c_int(device), _p(idb), c_int(rank), c_int(nranks)))
        else:
            self._check(self.L.sq_ctx_create(ctypes.byref(self._ctx), c_int(device)))
        self.n = 0

    def _check(self, rc):
        if rc != 0:
            raise ShenqiError("shenqi_b200 error %d: %s" % (rc, self.L.sq_last_error().decode()))

    def close(self):
        if self._ctx:
            self.L.sq_ctx_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters: set_gravshort_treepar / gravshort_set_softenings / gravpm_init_periodic
    def set_params(self, params):
        self.params = params
        self._check(self.L.sq_set_params(self._ctx, ctypes.byref(params)))

    # ---- particles
    def upload_particles(self, pos, mass, types=None):
        pos = np.ascontiguousarray(pos, np.float64); mass = np.ascontiguousarray(mass, np.float32)
        assert pos.shape == (len(mass), 3)
        types = None if types is None else np.ascontiguousarray(types, np.uint8)
        self.n = len(mass)
        self._check(self.L.sq_upload_particles_soa(self._ctx, _p(pos), _p(mass), None, _p(types), c_i64(self.n)))

    def upload_particle_data(self, parts):
        """parts: numpy array with PARTICLE_DTYPE (the reference's struct particle_data, read in place)."""
        assert parts.dtype.itemsize == 160 and parts.flags.c_contiguous
        self.n = len(parts)
        self._check(self.L.sq_upload_particles_aos(self._ctx, _p(parts), c_i64(self.n), None))

    def upload_particles_bigfile(self, path, ptype=1, first=0, count=-1):
        """particles of one type of a bigfile snapshot / IC file (petaio's "<type>/Position", "<type>/Mass" or Header.MassTable);
        returns Header.BoxSize."""
        box = c_dbl()
        self._check(self.L.sq_upload_particles_bigfile(self._ctx, path.encode(), c_int(ptype), c_i64(first), c_i64(count), ctypes.byref(box)))
        _, _, nrows, _ = bigfile_block_info(path, "%d/Position" % ptype)
        self.n = (nrows - first) if count < 0 else count
        return box.value

    def set_old_accel(self, fulltreegravaccel=None, gravpm=None):
        a = None if fulltreegravaccel is None else np.ascontiguousarray(fulltreegravaccel, np.float64)
        b = None if gravpm is None else np.ascontiguousarray(gravpm, np.float64)
        self._check(self.L.sq_set_old_accel(self._ctx, _p(a), _p(b)))

    def set_oldacc(self, oldacc):
        a = np.ascontiguousarray(oldacc, np.float64)
        assert len(a) == self.n
        self._check(self.L.sq_set_oldacc(self._ctx, _p(a)))

    # ---- PEANO + slots_gc_sorted
    def ph_sort(self):
        self._check(self.L.sq_ph_sort(self._ctx))

    def keys(self):
        k = np.zeros(self.n, np.uint64)
        self._check(self.L.sq_get_keys(self._ctx, _p(k)))
        return k

    def sorted_keys_perm(self):
        k = np.zeros(self.n, np.uint64); perm = np.zeros(self.n, np.int32)
        self._check(self.L.sq_get_sorted(self._ctx, _p(k), _p(perm)))
        return k, perm

    # ---- force_tree_full (active=None) / force_tree_active_moments
    def tree_build(self, active=None):
        info = TreeInfo()
        act = None if active is None else np.ascontiguousarray(active, np.int32)
        self._check(self.L.sq_tree_build(self._ctx, _p(act), c_i64(0 if act is None else len(act)), ctypes.byref(info)))
        self.tree_info = info
        return info

    def tree_nodes(self, firstnode=None, upload_order=False):
        """The tree as the reference's NODE[] image."""
        firstnode = self.n if firstnode is None else firstnode
        nodes = np.zeros(self.tree_info.numnodes, NODE_DTYPE)
        self._check(self.L.sq_tree_export_nodes(self._ctx, _p(nodes), c_i64(firstnode), c_int(1 if upload_order else 0)))
        return nodes, firstnode

    # ---- multi-rank tree under the reference's DomainDecomp
    def domain_set(self, dom):
        self.domain = dom
        if dom is None:
            self._check(self.L.sq_domain_set(self._ctx, None, c_int(0), None, c_int(0), None, c_int(0)))
        else:
            self._check(self.L.sq_domain_set(self._ctx, _p(dom.topnodes), c_int(len(dom.topnodes)), _p(dom.topleaves), c_int(len(dom.topleaves)),
                                             _p(dom.tasks), c_int(len(dom.tasks))))

    def domain_histogram(self, level, weights=None):
        """key histogram of the uploaded particles over the level's 8^level cells, summed over ranks; returns (count, cost)."""
        count = np.zeros(8 ** level, np.int64); cost = np.zeros(8 ** level, np.float64)
        w = None if weights is None else np.ascontiguousarray(weights, np.float32)
        self._check(self.L.sq_domain_histogram(self._ctx, c_int(level), _p(count), _p(cost), _p(w)))
        return count, cost

    def keys_device(self, pos_ptr, n, keys_ptr):
        """PEANO keys of device positions into a device uint64 array (queued on the context's stream)."""
        self._check(self.L.sq_keys_device(self._ctx, ctypes.c_void_p(pos_ptr), c_i64(n), ctypes.c_void_p(keys_ptr)))

    def domain_owner(self):
        owner = np.zeros(self.n, np.int32)
        self._check(self.L.sq_domain_owner(self._ctx, _p(owner)))
        return owner

    def domain_treenodes(self):
        tn = np.zeros(self.domain.ntopleaves, np.int32)
        self._check(self.L.sq_domain_get_treenodes(self._ctx, _p(tn)))
        return tn

    def tree_build_dist(self):
        """sq_tree_build_dist (collective): returns (TreeInfo, LetInfo)."""
        info, let = TreeInfo(), LetInfo()
        self._check(self.L.sq_tree_build_dist(self._ctx, ctypes.byref(info), ctypes.byref(let)))
        self.tree_info = info
        return info, let

    def toptree_export_count(self):
        a, b = c_i64(), c_i64()
        self._check(self.L.sq_toptree_export_count(self._ctx, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    # ---- grav_short_tree
    def grav_short_tree(self, active=None, accel=None, potential=None):
        """Returns (accel[N,3] x G, potential[N] or None, WalkStats). accel/potential in upload order;
        entries of inactive particles are left as passed in."""
        st = WalkStats()
        act = None if active is None else np.ascontiguousarray(active, np.int32)
        accel = np.zeros((self.n, 3)) if accel is None else accel
        potential = np.zeros(self.n) if potential is None else potential
        self._check(self.L.sq_tree_walk(self._ctx, _p(act), c_i64(0 if act is None else len(act)), _p(accel), _p(potential), ctypes.byref(st)))
        return accel, potential, st

    def grav_short_tree_device(self, active=None):
        st = WalkStats()
        act = None if active is None else np.ascontiguousarray(active, np.int32)
        self._check(self.L.sq_tree_walk_device(self._ctx, _p(act), c_i64(0 if act is None else len(act)), ctypes.byref(st)))
        return st

    def tree_build_device_list(self, active_ptr, nactive):
        """force_tree_active_moments with a DEVICE-resident active list (e.g. from build_active_sublist_device)."""
        info = TreeInfo()
        self._check(self.L.sq_tree_build(self._ctx, ctypes.c_void_p(active_ptr), c_i64(nactive), ctypes.byref(info)))
        self.tree_info = info
        return info

    def grav_short_tree_device_list(self, active_ptr, nactive):
        """grav_short_tree over a DEVICE-resident active list; results stay on the device (download_accel / device_ptr)."""
        st = WalkStats()
        self._check(self.L.sq_tree_walk_device(self._ctx, ctypes.c_void_p(active_ptr), c_i64(nactive), ctypes.byref(st)))
        return st

    def download_accel(self):
        accel = np.zeros((self.n, 3)); pot = np.zeros(self.n)
        self._check(self.L.sq_download_accel(self._ctx, _p(accel), _p(pot)))
        return accel, pot

    def download_accel_device(self, accel_ptr, pot_ptr=None):
        self._check(self.L.sq_download_accel_device(self._ctx, ctypes.c_void_p(accel_ptr), ctypes.c_void_p(pot_ptr)))

    def walk_particle_cost(self, out_ptr=None):
        """per-particle work estimate of the last walk (upload order): numpy array, or written to a device pointer."""
        if out_ptr is not None:
            self._check(self.L.sq_walk_particle_cost(self._ctx, ctypes.c_void_p(out_ptr)))
            return None
        cost = np.zeros(self.n, np.float32)
        self._check(self.L.sq_walk_particle_cost(self._ctx, _p(cost)))
        return cost

    def set_oldacc_device(self, oldacc_ptr):
        """oldacc[i] (upload order) from a DEVICE array; call before ph_sort (the sort gathers it)."""
        self._check(self.L.sq_set_oldacc(self._ctx, ctypes.c_void_p(oldacc_ptr)))

    # ---- gravpm_force
    def gravpm_force(self):
        gravpm = np.zeros((self.n, 3)); pot = np.zeros(self.n)
        self._check(self.L.sq_pm_force(self._ctx, _p(gravpm), _p(pot)))
        return gravpm, pot

    def gravpm_force_device(self):
        self._check(self.L.sq_pm_force_device(self._ctx))

    # ---- one PM step of run(): gravpm_force + force_tree_full + grav_short_tree (run.cpp:538-563), copies overlapped
    def gravity_step(self, pos, mass, fulltree=None, gravpm_old=None, accel=None, potential=None, gravpm=None, pm_potential=None, pm=True):
        """pos [N,3] f64, mass [N] f32 and the previous step's FullTreeGravAccel / GravPM [N,3] (or None) on the host
        (pinned buffers overlap best).  Returns (accel x G, tree potential, gravpm, pm potential, WalkStats)."""
        assert pos.dtype == np.float64 and pos.flags.c_contiguous and mass.dtype == np.float32 and mass.flags.c_contiguous
        n = self.n = len(mass)
        for a in (fulltree, gravpm_old):
            assert a is None or (a.dtype == np.float64 and a.flags.c_contiguous and a.shape == (n, 3))
        accel = np.zeros((n, 3)) if accel is None else accel
        potential = np.zeros(n) if potential is None else potential
        if pm:
            gravpm = np.zeros((n, 3)) if gravpm is None else gravpm
            pm_potential = np.zeros(n) if pm_potential is None else pm_potential
        else:
            gravpm = pm_potential = None
        st = WalkStats()
        self._check(self.L.sq_gravity_step(self._ctx, _p(pos), _p(mass), c_i64(n), _p(fulltree), _p(gravpm_old), _p(accel), _p(potential),
                                           _p(gravpm), _p(pm_potential), ctypes.byref(st)))
        return accel, potential, gravpm, pm_potential, st

    def pm_paint(self):
        self._check(self.L.sq_pm_paint(self._ctx))
        n = self.params.Nmesh
        mesh = np.zeros((n, n, n))
        self._check(self.L.sq_pm_get_mesh(self._ctx, _p(mesh)))
        return mesh

    def timings(self):
        t = Timings()
        self._check(self.L.sq_get_timings(self._ctx, ctypes.byref(t)))
        return t

    def reset_timings(self):
        self._check(self.L.sq_reset_timings(self._ctx))

    def measure_fp32_peak(self):
        t = c_dbl()
        self._check(self.L.sq_measure_fp32_peak(self._ctx, ctypes.byref(t)))
        return t.value

    def set_group_width(self, g):
        """Targets per warp in the walk (32, 16 or 8): tuning knob."""
        assert g in (8, 16, 32)
        ctypes.c_int.in_dll(self.L, "sq_walk_group_width").value = g

    # ---- multi-GPU entry points: raw device pointers (ints), see include/shenqi_b200.h and shenqi_b200/dist.py
    def set_stream(self, cuda_stream):
        self._check(self.L.sq_ctx_set_stream(self._ctx, ctypes.c_void_p(cuda_stream)))

    def device_ptr(self, which):
        p = ctypes.c_void_p(); n = c_i64()
        self._check(self.L.sq_device_ptr(self._ctx, c_int(which), ctypes.byref(p), ctypes.byref(n)))
        return p.value, n.value

    def upload_particles_device(self, pos_ptr, mass_ptr, n):
        self.n = n
        self._check(self.L.sq_upload_particles_device(self._ctx, ctypes.c_void_p(pos_ptr), ctypes.c_void_p(mass_ptr), c_i64(n)))

    def geo_mismatch(self):
        f = c_int()
        self._check(self.L.sq_get_geo_mismatch(self._ctx, ctypes.byref(f)))
        return f.value

    def set_sorted_device(self, keys_ptr, pos_ptr, mass_ptr, n, geo_mismatch=0):
        self.n = n
        self._check(self.L.sq_set_sorted_device(self._ctx, ctypes.c_void_p(keys_ptr), ctypes.c_void_p(pos_ptr), ctypes.c_void_p(mass_ptr), c_i64(n), c_int(geo_mismatch)))

    def set_oldacc_range_device(self, first, count, oldacc_ptr):
        self._check(self.L.sq_set_oldacc_range_device(self._ctx, c_i64(first), c_i64(count), ctypes.c_void_p(oldacc_ptr)))

    def tree_walk_range(self, first, count):
        st = WalkStats()
        self._check(self.L.sq_tree_walk_range(self._ctx, c_i64(first), c_i64(count), ctypes.byref(st)))
        return st

    def pm_slab_route(self, pos_ptr, mass_ptr, n, nranks, send4_ptr, slot_ptr):
        counts = (c_i64 * nranks)()
        self._check(self.L.sq_pm_slab_route(self._ctx, ctypes.c_void_p(pos_ptr), ctypes.c_void_p(mass_ptr), c_i64(n), c_int(nranks),
                                            ctypes.c_void_p(send4_ptr), ctypes.c_void_p(slot_ptr), counts))
        return [int(x) for x in counts]

    def pm_slab_unroute(self, back4_ptr, slot_ptr, n, out4_ptr):
        self._check(self.L.sq_pm_slab_unroute(self._ctx, ctypes.c_void_p(back4_ptr), ctypes.c_void_p(slot_ptr), c_i64(n), ctypes.c_void_p(out4_ptr)))

    def pm_slab_paint(self, part4_ptr, m, rank, nranks, slab_ptr):
        self._check(self.L.sq_pm_slab_paint(self._ctx, ctypes.c_void_p(part4_ptr), c_i64(m), c_int(rank), c_int(nranks), ctypes.c_void_p(slab_ptr)))

    def pm_slab_fft_fwd(self, slab_ptr, nranks, work_ptr, packed_ptr):
        self._check(self.L.sq_pm_slab_fft_fwd(self._ctx, ctypes.c_void_p(slab_ptr), c_int(nranks), ctypes.c_void_p(work_ptr), ctypes.c_void_p(packed_ptr)))

    def pm_slab_fft_x(self, data_ptr, nranks, inverse):
        self._check(self.L.sq_pm_slab_fft_x(self._ctx, ctypes.c_void_p(data_ptr), c_int(nranks), c_int(1 if inverse else 0)))

    def pm_slab_transfer(self, rho_ptr, rank, nranks, out4_ptr):
        self._check(self.L.sq_pm_slab_transfer(self._ctx, ctypes.c_void_p(rho_ptr), c_int(rank), c_int(nranks), ctypes.c_void_p(out4_ptr)))

    def pm_slab_transfer_phi(self, rho_ptr, rank, nranks, phik_ptr):
        self._check(self.L.sq_pm_slab_transfer_phi(self._ctx, ctypes.c_void_p(rho_ptr), c_int(rank), c_int(nranks), ctypes.c_void_p(phik_ptr)))

    def pm_slab_gradient(self, phi_ptr, lo2_ptr, hi3_ptr, nranks, ext_ptr, fx_ptr, fy_ptr, fz_ptr):
        self._check(self.L.sq_pm_slab_gradient(self._ctx, ctypes.c_void_p(phi_ptr), ctypes.c_void_p(lo2_ptr), ctypes.c_void_p(hi3_ptr), c_int(nranks),
                                               ctypes.c_void_p(ext_ptr), ctypes.c_void_p(fx_ptr), ctypes.c_void_p(fy_ptr), ctypes.c_void_p(fz_ptr)))

    def pm_slab_fft_inv(self, recv_ptr, nranks, work_ptr, slab_ptr):
        self._check(self.L.sq_pm_slab_fft_inv(self._ctx, ctypes.c_void_p(recv_ptr), c_int(nranks), ctypes.c_void_p(work_ptr), ctypes.c_void_p(slab_ptr)))

    def pm_slab_readout(self, part4_ptr, m, rank, nranks, slab_ptrs, out4_ptr):
        self._check(self.L.sq_pm_slab_readout(self._ctx, ctypes.c_void_p(part4_ptr), c_i64(m), c_int(rank), c_int(nranks),
                                              *[ctypes.c_void_p(p) for p in slab_ptrs], ctypes.c_void_p(out4_ptr)))

    # ---- drift_all_particles / apply_hierarchical_grav_kick / build_active_sublist on device arrays
    def drift_device(self, pos_ptr, vel_ptr, n, ddrift, random_shift, BoxSize, skip_ptr=None):
        """drift_all_particles (libgadget/drift.cpp:81-99): pos, vel = device double[n][3]; skip = device uint8[n] or None."""
        sh = (c_dbl * 3)(*[float(x) for x in random_shift]) if random_shift is not None else None
        self._check(self.L.sq_drift_device(self._ctx, ctypes.c_void_p(pos_ptr), ctypes.c_void_p(vel_ptr), ctypes.c_void_p(skip_ptr),
                                           c_i64(n), c_dbl(ddrift), sh, c_dbl(BoxSize)))

    def kick_device(self, vel_ptr, accel_ptr, nactive, gravkick, active_ptr=None, skip_ptr=None):
        """the kick loop of apply_hierarchical_grav_kick (libgadget/timestep.cpp:270-279, 962-968)."""
        self._check(self.L.sq_kick_device(self._ctx, ctypes.c_void_p(vel_ptr), ctypes.c_void_p(accel_ptr), ctypes.c_void_p(active_ptr),
                                          ctypes.c_void_p(skip_ptr), c_i64(nactive), c_dbl(gravkick)))

    def build_active_sublist_device(self, nactive, timebin_ptr, maxtimebin, Ti_Current, sublist_ptr, active_ptr=None, skip_ptr=None):
        """build_active_sublist (libgadget/timestep.cpp:1371-1397); returns NumActiveParticle of the sub-list."""
        nsub = c_i64(0)
        self._check(self.L.sq_active_sublist_device(self._ctx, ctypes.c_void_p(active_ptr), c_i64(nactive), ctypes.c_void_p(timebin_ptr),
                                                    ctypes.c_void_p(skip_ptr), c_int(maxtimebin), c_i64(Ti_Current),
                                                    ctypes.c_void_p(sublist_ptr), ctypes.byref(nsub)))
        return int(nsub.value)

    def gravity_timebins_device(self, par, accel_ptr, gravpm_ptr, nactive, timebin_ptr, active_ptr=None, skip_ptr=None):
        """the bin-assignment loop of hierarchical_gravity_accelerations (libgadget/timestep.cpp:351-377):
        returns (timebincounts[47] of this rank, number of bad time steps)."""
        counts = (c_i64 * (TIMEBINS + 1))()
        nbad = c_i64(0)
        self._check(self.L.sq_gravity_timebins_device(self._ctx, ctypes.byref(par), ctypes.c_void_p(accel_ptr), ctypes.c_void_p(gravpm_ptr),
                                                      ctypes.c_void_p(active_ptr), ctypes.c_void_p(skip_ptr), c_i64(nactive),
                                                      ctypes.c_void_p(timebin_ptr), counts, ctypes.byref(nbad)))
        return [int(x) for x in counts], int(nbad.value)

    def timebin_pushdown_device(self, nactive, timebin_ptr, push_down_bin, active_ptr=None):
        """libgadget/timestep.cpp:399-407."""
        self._check(self.L.sq_timebin_pushdown_device(self._ctx, ctypes.c_void_p(active_ptr), c_i64(nactive), ctypes.c_void_p(timebin_ptr), c_int(push_down_bin)))


def pushdown_bin(counts, largest_active, isPM):
    """Host side of libgadget/timestep.cpp:379-398 on the (all-reduced) timebincounts: the largest occupied bin, then on PM
    steps the push-down rule; returns (push_down_bin, counts after the merges).  0 is the reference's endrun(77)."""
    counts = list(counts)
    for ti in range(largest_active, 0, -1):
        if counts[ti] > 0:
            largest_active = ti
            break
    push = largest_active
    if isPM:
        for ti in range(largest_active, 0, -1):
            if counts[ti] // 3 > counts[ti - 1]:
                break
            push = ti - 1
            counts[ti - 1] += counts[ti]
    return push, counts
