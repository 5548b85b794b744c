"""CPU tests (no GPU) of the host logic of the domain-decomposed tree path: the domain builder of the C ABI
(sq_domain_from_histogram is pure host code: libgadget/domain.cpp's top-tree refinement and balanced leaf assignment),
the topnode_data / topleaf_data / task_data layouts, domain_get_topleaf, and - with world_size 2 over gloo - the flow the
multi-GPU bench follows: every rank histograms its own particles, the histograms are all-reduced, every rank derives
the SAME domain, and the owners partition the particle set."""
import os
import socket
import sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE); sys.path.insert(0, os.path.dirname(HERE))
from helpers import o_keys, random_clustered, ref


def _hist(keys, level):
    return np.bincount((keys >> np.uint64(63 - 3 * level)).astype(np.int64), minlength=8 ** level)


def _check_domain(D, nranks, keys, level):
    T, L, K = D.topnodes, D.topleaves, D.tasks
    assert T["Shift"][0] == 63 and T["StartKey"][0] == 0                       # root = the whole key space (domain.cpp:1076)
    inner = np.flatnonzero(T["Daughter"] >= 0)
    for i in inner:                                                              # domain.cpp:849-868: 8 daughters, key order
        d = T["Daughter"][i]
        assert np.all(T["Shift"][d:d + 8] == T["Shift"][i] - 3)
        assert np.array_equal(T["StartKey"][d:d + 8], T["StartKey"][i] + (np.arange(8, dtype=np.uint64) << np.uint64(T["Shift"][i] - 3)))
    leaves = np.flatnonzero(T["Daughter"] < 0)
    assert len(leaves) == len(L) and sorted(T["Leaf"][leaves]) == list(range(len(L)))
    assert np.array_equal(L["topnode"][T["Leaf"][leaves]], leaves)
    # leaves in key order, cut into nranks contiguous segments (domain_assign_topleaves_balanced)
    sk = T["StartKey"][L["topnode"]].astype(np.float64)
    assert np.all(np.diff(sk) > 0)
    assert K["StartLeaf"][0] == 0 and K["EndLeaf"][-1] == len(L) and np.array_equal(K["StartLeaf"][1:], K["EndLeaf"][:-1])
    for t in range(nranks):
        assert K["EndLeaf"][t] > K["StartLeaf"][t]
        assert np.all(L["Task"][K["StartLeaf"][t]:K["EndLeaf"][t]] == t)
    # every key falls into exactly the leaf whose key range holds it
    leaf = D.topleaf_of(keys)
    tn = L["topnode"][leaf]
    assert np.all(keys >= T["StartKey"][tn]) and np.all((keys - T["StartKey"][tn]) >> T["Shift"][tn].astype(np.uint64) == 0)
    return L["Task"][leaf]


@pytest.mark.parametrize("nranks,level,over", [(1, 2, 2), (2, 3, 4), (8, 3, 4), (8, 4, 4), (5, 3, 8)])
def test_domain_from_histogram(nranks, level, over):
    import shenqi_b200 as sq
    box = 8.0
    pos = random_clustered(100000, box, 7)
    keys = o_keys(pos, box)
    cnt = _hist(keys, level)
    D = sq.domain_from_histogram(cnt, nranks, over)
    owner = _check_domain(D, nranks, keys, level)
    load = np.bincount(owner, minlength=nranks)
    assert load.sum() == len(pos) and load.min() > 0
    # no leaf above the refinement limit unless it sits at the histogram's own resolution
    lim = len(pos) / (over * nranks)
    T = D.topnodes[D.topleaves["topnode"]]
    per_leaf = np.bincount(D.topleaf_of(keys), minlength=D.ntopleaves)
    assert np.all((per_leaf <= lim) | (T["Shift"] == 63 - 3 * level))
    # cost-weighted: segments of equal cost, not equal count
    cost = cnt.astype(np.float64) ** 1.5
    Dc = sq.domain_from_histogram(cnt, nranks, over, cost=cost)
    oc = _check_domain(Dc, nranks, keys, level)
    cell = (keys >> np.uint64(63 - 3 * level)).astype(np.int64)
    w = (cost / np.maximum(cnt, 1))[cell]
    if nranks > 1:
        # a greedy cut is never further from the mean than the heaviest leaf
        per_rank = np.bincount(oc, weights=w, minlength=nranks)
        heaviest = np.bincount(Dc.topleaf_of(keys), weights=w, minlength=Dc.ntopleaves).max()
        assert np.abs(per_rank - w.sum() / nranks).max() <= heaviest * (1 + 1e-9)


def test_domain_struct_layouts_match_reference():
    import shenqi_b200 as sq
    assert sq.TOPNODE_DTYPE.itemsize == 24 and sq.TOPLEAF_DTYPE.itemsize == 12 and sq.TASK_DTYPE.itemsize == 8     # domain.h:12-31
    if ref() is not None:
        assert ref().ref_sizeof_topnode() == 24 and ref().ref_sizeof_topleaf() == 12


def test_domain_errors():
    import shenqi_b200 as sq
    with pytest.raises(sq.ShenqiError):
        sq.domain_from_histogram(np.array([5], np.int64), 2, 1)          # one cell, two ranks: "Number of Topleaves is less than NTask"


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import shenqi_b200 as sq
        box, level = 8.0, 3
        pos = random_clustered(60000, box, 11)
        keys = o_keys(pos, box)
        mine = np.arange(rank, len(pos), world)                         # before the exchange: particles anywhere
        h = torch.from_numpy(_hist(keys[mine], level))
        dist.all_reduce(h)                                               # MPI_Allreduce of the local top-tree counts (domain.cpp:1160-1230)
        D = sq.domain_from_histogram(h.numpy(), world, 4)
        owner = D.topleaves["Task"][D.topleaf_of(keys[mine])]
        # what the particle exchange would move: counts to every rank, all-to-all of the counts must be consistent
        send = torch.from_numpy(np.bincount(owner, minlength=world))
        recv = [torch.zeros_like(send) for _ in range(world)]
        dist.all_gather(recv, send)
        q.put((rank, [D.topnodes[f].tolist() for f in D.topnodes.dtype.names], D.topleaves.tobytes(), D.tasks.tobytes(), h.numpy().sum(), torch.stack(recv).numpy()))
    finally:
        dist.destroy_process_group()


def test_world2_gloo_same_domain_on_every_rank():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = {}
    for _ in range(world):
        r, a, b, c, tot, table = q.get(timeout=300)
        out[r] = (a, b, c, tot, table)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert out[0][:3] == out[1][:3]                                      # the same DomainDecomp on both ranks, field for field
    assert out[0][3] == out[1][3] == 60000
    table = out[0][4]                                                    # table[s][d] = particles rank s sends to rank d
    assert table.sum() == 60000 and np.all(table.sum(axis=0) > 0)


def test_bench_topleaf_lookup_equals_domain_get_topleaf():
    """bench.py's device-side owner lookup (torch ops) is domain_get_topleaf (domain.h:68-75): checked on CPU tensors."""
    import sys
    sys.path.insert(0, os.path.dirname(HERE))
    import bench
    import shenqi_b200 as sq
    box = 8.0
    pos = random_clustered(50000, box, 3)
    keys = o_keys(pos, box)
    D = sq.domain_from_histogram(_hist(keys, 4), 8, 4)
    got = bench.torch_topleaf(D, torch.from_numpy(keys.astype(np.int64))).numpy()
    assert np.array_equal(got, D.topleaf_of(keys))


# ------------------------------------------------------------------------------------------ pinned to the reference's domain.cpp
def _has_ref_domain():
    R = ref()
    return R is not None and hasattr(R, "ref_domain_assign_balanced")


needs_ref_domain = pytest.mark.skipif(not _has_ref_domain(), reason="oracle/_ref was built without the reference's domain.cpp")


def _ref_assign(cost, ntask, nspt, rng):
    """domain_assign_topleaves_balanced + domain_set_task_leafs of the reference itself (oracle/ref_domain.cpp) on leaves
    handed over in scrambled order; returns (task of every leaf in key order, tasks table)."""
    import ctypes
    import shenqi_b200 as sq
    from helpers import ptr, quiet_stdout
    nl = len(cost)
    tn = np.zeros(nl + 1, sq.TOPNODE_DTYPE)                     # a root and one top node per leaf, StartKey ascending
    tn["StartKey"][1:] = np.arange(nl, dtype=np.uint64) * np.uint64(8)
    tn["Daughter"] = -1; tn["Shift"] = 3
    perm = rng.permutation(nl)
    leaf_topnode = (1 + perm).astype(np.int32)
    task = np.zeros(nl, np.int32); topn = np.zeros(nl, np.int32); tasks = np.zeros(2 * ntask, np.int32)
    with quiet_stdout():
        ref().ref_domain_assign_balanced(ctypes.c_int(ntask), ctypes.c_int(nspt), ctypes.c_int(nl + 1), ptr(tn), ctypes.c_int(nl), ptr(leaf_topnode),
                                         ptr(np.ascontiguousarray(cost[perm])), ptr(task), ptr(topn), ptr(tasks))
    by_key = np.zeros(nl, np.int32)
    by_key[topn - 1] = task                                     # the reference returns the leaves sorted by (Task, Key)
    return by_key, tasks.reshape(ntask, 2)


@needs_ref_domain
def test_assign_balanced_equals_reference_domain_cpp():
    """sq_domain_assign_balanced against the reference's OWN domain_assign_topleaves_balanced (domain.cpp:620-778): the
    same task for EVERY leaf, for 300 random (task count, segments per task, leaf count, cost distribution) draws -
    uniform, heavy-tailed (a few leaves hold most of the cost), constant, and mostly-zero costs."""
    import shenqi_b200 as sq
    rng = np.random.default_rng(0)
    for trial in range(300):
        ntask = int(rng.integers(1, 17)); nspt = int(rng.choice([1, 1, 1, 2, 4]))
        nl = int(rng.integers(ntask * nspt, ntask * nspt * 12 + 2))
        kind = trial % 4
        if kind == 0:
            cost = rng.integers(0, 1000, nl)
        elif kind == 1:
            cost = (rng.pareto(1.2, nl) * 100).astype(np.int64)
        elif kind == 2:
            cost = np.full(nl, 7)
        else:
            cost = rng.integers(0, 3, nl) * rng.integers(1, 10 ** 6, nl)
        cost = cost.astype(np.int64)
        want, tasks = _ref_assign(cost, ntask, nspt, rng)
        got = sq.domain_assign_balanced(cost, ntask, nspt)
        assert np.array_equal(got, want), (trial, ntask, nspt, nl, kind)
        if nspt == 1:                                           # one segment per task: contiguous, every task served
            assert np.all(np.diff(got) >= 0) and len(np.unique(got)) == ntask
            for t in range(ntask):
                mine = np.flatnonzero(got == t)
                assert (tasks[t, 0], tasks[t, 1]) == (mine[0], mine[-1] + 1)


def _ref_decompose(pos, box, over):
    """domain_decompose_full (domain.cpp:174-277) of the reference on one rank; returns (Domain, Part[].TopLeaf in the
    reference's sorted order, sorted positions)."""
    import ctypes
    import shenqi_b200 as sq
    from helpers import ptr, quiet_stdout, r_load, r_sorted_particles
    R = ref()
    n = len(pos)
    tn = np.zeros(200000, sq.TOPNODE_DTYPE); tl = np.zeros(200000, sq.TOPLEAF_DTYPE); ntl = ctypes.c_int()
    with quiet_stdout():
        r_load(pos, np.ones(n, np.float32), box, ids=np.arange(1, n + 1, dtype=np.uint64))
        nn = R.ref_domain_decompose_full(ctypes.c_int(over), ctypes.c_double(0.5), ptr(tn), ctypes.c_int(len(tn)), ptr(tl), ctypes.c_int(len(tl)),
                                         ctypes.byref(ntl))
    D = sq.Domain(tn[:nn].copy(), tl[:ntl.value].copy(), np.array([(0, ntl.value)], sq.TASK_DTYPE))
    topleaf = np.zeros(n, np.int32)
    R.ref_domain_get_topleaf(ptr(topleaf))
    spos, _, _ = r_sorted_particles()
    return D, topleaf, spos


@needs_ref_domain
@pytest.mark.parametrize("over,seed", [(4, 21), (16, 22), (64, 23)])
def test_domain_of_reference_domain_decompose_full_is_accepted(over, seed):
    """a DomainDecomp the reference ITSELF builds (domain_decompose_full on clustered particles: subsampled top tree,
    refinement, truncation, leaves of unequal depth) passes the checks sq_domain_set applies (sq_domain_validate), and
    the host-side domain_get_topleaf restatement (Domain.topleaf_of) gives Part[i].TopLeaf for every particle."""
    box = 8.0
    pos = random_clustered(60000, box, seed)
    D, topleaf, spos = _ref_decompose(pos, box, over)
    assert D.ntopleaves >= over
    maxlevel = D.validate()
    lv = (63 - D.topnodes["Shift"]) // 3
    assert maxlevel == lv.max() and len(np.unique(lv[D.topnodes["Daughter"] < 0])) > 1          # leaves of unequal depth
    assert np.array_equal(D.topleaf_of(o_keys(spos, box)), topleaf)
    # slots_gc_sorted at the end of domain_decompose_full leaves the particles in key order: leaves are then contiguous
    assert np.all(np.diff(topleaf) >= 0)


def test_domain_validate_refuses_broken_domains():
    import shenqi_b200 as sq
    keys = o_keys(random_clustered(20000, 8.0, 1), 8.0)
    D = sq.domain_from_histogram(_hist(keys, 3), 4, 4)
    assert D.validate() >= 1
    for breaker in ("startkey", "leaf", "task", "range"):
        B = sq.Domain(D.topnodes.copy(), D.topleaves.copy(), D.tasks.copy())
        if breaker == "startkey":
            B.topnodes["StartKey"][B.topnodes["Daughter"][0] + 3] += np.uint64(1)
        elif breaker == "leaf":
            B.topnodes["Leaf"][np.flatnonzero(B.topnodes["Daughter"] < 0)[0]] = B.ntopleaves
        elif breaker == "task":
            B.topleaves["Task"][0] = 4
        else:
            B.tasks["EndLeaf"][0] += 1
        with pytest.raises(sq.ShenqiError, match="inconsistent DomainDecomp"):
            B.validate()


@pytest.mark.parametrize("nranks,level,over", [(2, 3, 4), (8, 3, 4), (8, 4, 8), (5, 3, 8)])
def test_domain_from_histogram_with_the_reference_assignment(nranks, level, over):
    """assign="reference": the same top tree, leaves handed to tasks by the reference's greedy rule; a valid domain whose
    per-task load stays within one (heaviest) leaf of the mean, with cost weights as well."""
    import shenqi_b200 as sq
    box = 8.0
    pos = random_clustered(100000, box, 7)
    keys = o_keys(pos, box)
    cnt = _hist(keys, level)
    for cost in (None, cnt * (1.0 + (np.arange(len(cnt)) % 5))):
        D = sq.domain_from_histogram(cnt, nranks, over, cost=cost, assign="reference")
        D.validate()
        owner = _check_domain(D, nranks, keys, level)
        w = np.ones(len(keys)) if cost is None else (1.0 + ((keys >> np.uint64(63 - 3 * level)).astype(np.int64) % 5))
        load = np.bincount(owner, weights=w, minlength=nranks)
        leafload = np.bincount(D.topleaf_of(keys), weights=w, minlength=D.ntopleaves)
        assert load.min() > 0 and np.abs(load - load.mean()).max() <= leafload.max() + 1e-9
