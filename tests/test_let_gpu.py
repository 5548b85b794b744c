"""Multi-rank force tree on real devices (SURVEY rows a5 top nodes, a7, a14, a15, e): the tree under the reference's
DomainDecomp, top-leaf moment exchange, import of the reachable remote sub-trees, all through the C ABI.

  * one rank, multi-leaf domain: the forced top-level nodes against the reference's OWN force_tree_full with the same
    DomainDecomp (oracle/_ref): canonical topology incl. TopLevel / InternalTopLevel flags, TopLeaves[].treenode cells,
    moments, interaction counts and accelerations;
  * R = 2 and 4 ranks as host threads on ONE GPU over the library's in-process test transport (NCCL refuses two ranks on
    one device): the summed interaction count equals the one-rank count with the same domain EXACTLY, no pseudo node is
    opened without an import below it, accelerations agree to the FP32 pair-sum rounding (groups of 32 start elsewhere);
  * R = 2 over NCCL (one process per GPU) when the box has two GPUs: same assertions."""
import os
import sys
import threading
import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE); sys.path.insert(0, os.path.dirname(HERE))
from helpers import (ref, ptr, c_i64, c_dbl, c_int, quiet_stdout, canonical_tree, rel_err, default_params, r_load, r_tree, r_walk, r_counts,
                     r_sorted_particles, random_clustered, o_keys)

pytestmark = pytest.mark.gpu


def _face_coordinate(box, depth, rng):
    """a coordinate exactly on a face between two level-(depth + 1) cells: the centre of a level-`depth` cell, by the
    reference's own recurrence centre +- len / 4 (forcetree.cpp:277-283, 304-319)"""
    c, q = box / 2., 0.25 * (box * 1.001)
    for _ in range(depth):
        c = c + q if rng.random() < 0.5 else c - q
        q *= 0.5
    return c


def _problem(n=60000, seed=11, nmesh=48, onface=0):
    import shenqi_b200 as sq
    box = 8.0
    pos = random_clustered(n, box, seed)
    if onface:
        # particles exactly on faces of cells BELOW the domain's top levels: key digits and geometric octants disagree there,
        # the tree is ordered by geometric keys (the reference descends by `Pos > centre`)
        rng = np.random.default_rng(seed + 1)
        for i in rng.choice(n, onface, replace=False):
            pos[i, rng.integers(3)] = _face_coordinate(box, int(rng.integers(6, 14)), rng)
    mass = (1.0 + 0.25 * (np.arange(n) % 3)).astype(np.float32)
    P = sq.Params.defaults(box, nmesh, (box / n ** (1 / 3.)) / 30., TreeUseBH=1, TreeRcut=6.0)
    return box, pos, mass, P


def _domain(pos, box, nranks, level=3, over=4):
    import shenqi_b200 as sq
    keys = o_keys(pos, box)
    cnt = np.bincount((keys >> np.uint64(63 - 3 * level)).astype(np.int64), minlength=8 ** level)
    return sq.domain_from_histogram(cnt, nranks, over), keys


def _single_rank_with_domain(pos, mass, P, dom, oldacc_from=None):
    """one context, all particles, the domain's forced top nodes: the reference for the multi-rank runs"""
    import shenqi_b200 as sq
    import copy
    d1 = sq.Domain(dom.topnodes.copy(), dom.topleaves.copy(), np.array([(0, dom.ntopleaves)], sq.TASK_DTYPE))
    d1.topleaves["Task"] = 0
    c = sq.Context(0)
    try:
        c.set_params(P)
        c.upload_particles(pos, mass)
        if oldacc_from is not None:
            c.set_old_accel(oldacc_from, None)
        c.ph_sort()
        c.domain_set(d1)
        info, let = c.tree_build_dist()
        acc, pot, st = c.grav_short_tree()
        nodes, first = c.tree_nodes(firstnode=len(pos), upload_order=True)
        tn = c.domain_treenodes()
    finally:
        c.close()
    return acc, pot, st, nodes, first, tn, info


@pytest.mark.skipif(ref() is None, reason="reference build (oracle/_ref) not present")
@pytest.mark.parametrize("level,over", [(1, 2), (2, 4), (3, 8)])
def test_domain_top_nodes_against_reference_build(level, over):
    import shenqi_b200 as sq
    box, pos, mass, P = _problem(40000, 3)
    n = len(pos)
    dom, keys = _domain(pos, box, 1, level, over)
    assert dom.ntopleaves >= (8 if level == 1 else 15)
    acc, pot, st, nodes, first, tn, info = _single_rank_with_domain(pos, mass, P, dom)
    ids = np.arange(1, n + 1, dtype=np.uint64)
    R = ref()
    opar = default_params(box, P.Nmesh, n, TreeUseBH=1, TreeRcut=P.TreeRcut, GravitySoftening=P.GravitySoftening, rho0=P.rho0)
    with quiet_stdout():
        r_load(pos, mass, box, ids=ids)
        R.ref_sort_particles()
        _, _, rids = r_sorted_particles()
        R.ref_set_domain(ptr(dom.topnodes), c_int(len(dom.topnodes)), ptr(dom.topleaves), c_int(dom.ntopleaves))
        rnodes, rfirst, _ = r_tree(0.9)
        rtn = np.zeros(dom.ntopleaves, np.int32)
        R.ref_domain_treenodes(ptr(rtn))
        ra, _ = r_walk(opar, box / n ** (1 / 3.))
        _, rtot = r_counts(opar, box / n ** (1 / 3.))
        R.ref_tree_free()
        R.ref_set_domain(None, c_int(0), None, c_int(0))
    inv = (rids - 1).astype(np.int64)
    cells, parts, moms = canonical_tree(nodes, first, ids, box)
    rcells, rparts, rmoms = canonical_tree(rnodes, rfirst, rids, box)
    # columns: level, ix, iy, iz, ChildType, TopLevel, InternalTopLevel, particles in the leaf
    assert cells[:, 5].sum() == len(dom.topnodes) and cells[:, 6].sum() == len(dom.topnodes) - dom.ntopleaves
    assert np.array_equal(cells, rcells) and np.array_equal(parts, rparts)
    assert (np.abs(moms - rmoms) / np.abs(rmoms).max(axis=0)).max() < 1e-12
    # TopLeaves[].treenode points at the same cells
    def cell_of(nd, idx):
        rootlen = nd["len"][0]; corner = nd["center"][0] - rootlen / 2
        lv = np.rint(np.log2(rootlen / nd["len"][idx])).astype(np.int64)
        return np.concatenate([lv[:, None], np.floor((nd["center"][idx] - corner) / nd["len"][idx][:, None]).astype(np.int64)], 1)
    assert np.array_equal(cell_of(nodes, tn), cell_of(rnodes, rtn))
    assert st.interactions == int(rtot), (st.interactions, int(rtot))
    e = rel_err(acc[inv], ra)
    assert np.median(e) <= 1e-5 and np.percentile(e, 99.9) <= 1e-4, (np.median(e), np.percentile(e, 99.9))
    assert st.pseudo_opened == 0


def _rank_body(rank, R, nccl_id, device, pos, mass, P, dom, keys, oldacc_full, out, relative):
    import shenqi_b200 as sq
    try:
        owner = dom.topleaves["Task"][dom.topleaf_of(keys)]
        mine = np.flatnonzero(owner == rank)
        np.random.default_rng(100 + rank).shuffle(mine)
        c = sq.Context(device, nccl_id=nccl_id, rank=rank, nranks=R)
        try:
            Pr = sq.Params.defaults(P.BoxSize, P.Nmesh, P.GravitySoftening, TreeUseBH=0 if relative else 1, TreeRcut=P.TreeRcut)
            c.set_params(Pr)
            c.upload_particles(np.ascontiguousarray(pos[mine]), np.ascontiguousarray(mass[mine]))
            # the collective histogram gives every rank the same domain the test built on the host
            cnt, _ = c.domain_histogram(3)
            if relative:
                c.set_old_accel(np.ascontiguousarray(oldacc_full[mine]), None)
            c.ph_sort()
            c.domain_set(dom)
            assert np.all(c.domain_owner() == rank)
            info, let = c.tree_build_dist()
            nexp, npart = c.toptree_export_count()
            acc, pot, st = c.grav_short_tree()
            out[rank] = dict(mine=mine, acc=acc, pot=pot, inter=st.interactions, pseudo=st.pseudo_opened, hist=cnt,
                             let={k: getattr(let, k) for k, _ in let._fields_}, nexp=nexp, npart=npart, nodes=info.numnodes)
        finally:
            c.close()
    except Exception as e:          # noqa
        import traceback
        out[rank] = dict(error=traceback.format_exc())


def _check_ranks(out, R, pos, ref_acc, ref_pot, ref_inter, keys):
    for r in range(R):
        assert "error" not in out[r], out[r].get("error")
    assert sum(out[r]["inter"] for r in range(R)) == ref_inter, ([out[r]["inter"] for r in range(R)], ref_inter)
    cnt = np.bincount((keys >> np.uint64(63 - 9)).astype(np.int64), minlength=512)
    for r in range(R):
        o = out[r]
        assert o["pseudo"] == 0
        assert np.array_equal(o["hist"], cnt)
        m = o["mine"]
        e = rel_err(o["acc"], ref_acc[m])
        assert np.median(e) <= 1e-6 and np.percentile(e, 99.9) <= 1e-4, (r, np.median(e), np.percentile(e, 99.9))
        ep = np.abs(o["pot"] - ref_pot[m]) / np.abs(ref_pot[m])
        assert np.median(ep) <= 1e-6 and np.percentile(ep, 99.9) <= 1e-4, (r, np.median(ep), ep.max())
        L = o["let"]
        # memory is O(N/P + boundary), not O(N): the import is a fraction of what the other ranks hold
        assert 0 < L["imported_particles"] < len(pos) - len(m)
        assert L["imported_nodes"] > 0 and L["exported_nodes"] > 0


@pytest.mark.parametrize("R,relative,onface", [(2, False, 0), (2, True, 0), (4, True, 0), (2, True, 40)])
def test_let_ranks_as_threads_on_one_gpu(R, relative, onface):
    import shenqi_b200 as sq
    box, pos, mass, P = _problem(onface=onface)
    dom, keys = _domain(pos, box, R, 3, 4)
    # |a_old| for the relative criterion from a Barnes-Hut pass (as the reference bootstraps)
    a_bh, _, _, _, _, _, _ = _single_rank_with_domain(pos, mass, P, dom)
    P1 = sq.Params.defaults(P.BoxSize, P.Nmesh, P.GravitySoftening, TreeUseBH=0 if relative else 1, TreeRcut=P.TreeRcut)
    ref_acc, ref_pot, st, _, _, _, _ = _single_rank_with_domain(pos, mass, P1, dom, oldacc_from=a_bh if relative else None)
    nid = sq.comm_loopback_id(1000 + 10 * R + int(relative) + 100 * int(onface > 0))
    out = {}
    th = [threading.Thread(target=_rank_body, args=(r, R, nid, 0, pos, mass, P, dom, keys, a_bh, out, relative), daemon=True) for r in range(R)]
    for t in th:
        t.start()
    for t in th:
        t.join(600)
        assert not t.is_alive()
    _check_ranks(out, R, pos, ref_acc, ref_pot, st.interactions, keys)
    # the export decision of the reference (toptree_visit) is made for the same particles: some, not all, export
    assert 0 < sum(out[r]["npart"] for r in range(R)) < len(pos)


def _nccl_worker(rank, R, q_id, q_out, relative):
    import shenqi_b200 as sq
    torch.cuda.set_device(rank)
    box, pos, mass, P = _problem()
    dom, keys = _domain(pos, box, R, 3, 4)
    a_bh = None
    if relative:
        a_bh, _, _, _, _, _, _ = _single_rank_with_domain(pos, mass, P, dom)
    if rank == 0:
        nid = sq.comm_unique_id()
        for _ in range(R - 1):
            q_id.put(nid.tobytes())
    else:
        nid = np.frombuffer(q_id.get(timeout=300), np.uint8).copy()
    out = {}
    _rank_body(rank, R, nid, rank, pos, mass, P, dom, keys, a_bh, out, relative)
    q_out.put((rank, out[rank]))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("relative", [False, True])
def test_let_world2_nccl(relative):
    import torch.multiprocessing as mp
    import shenqi_b200 as sq
    R = 2
    box, pos, mass, P = _problem()
    dom, keys = _domain(pos, box, R, 3, 4)
    a_bh, _, _, _, _, _, _ = _single_rank_with_domain(pos, mass, P, dom)
    P1 = sq.Params.defaults(P.BoxSize, P.Nmesh, P.GravitySoftening, TreeUseBH=0 if relative else 1, TreeRcut=P.TreeRcut)
    ref_acc, ref_pot, st, _, _, _, _ = _single_rank_with_domain(pos, mass, P1, dom, oldacc_from=a_bh if relative else None)
    ctx = mp.get_context("spawn")
    q_id, q_out = ctx.Queue(), ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, R, q_id, q_out, relative)) for r in range(R)]
    for p in procs:
        p.start()
    out = {}
    for _ in range(R):
        r, res = q_out.get(timeout=900)
        out[r] = res
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    _check_ranks(out, R, pos, ref_acc, ref_pot, st.interactions, keys)
