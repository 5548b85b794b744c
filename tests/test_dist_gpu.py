"""Multi-GPU path on real devices through the C ABI (shenqi_b200.dist.CudaBackend): world 1 on one B200 against the
single-context path, and world 2 over NCCL when the box has two GPUs."""
import os
import socket
import sys
import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE); sys.path.insert(0, os.path.dirname(HERE))
pytestmark = pytest.mark.gpu


def _problem(n=40000, seed=5):
    from helpers import random_clustered
    import shenqi_b200 as sq
    box, nmesh = 8.0, 48
    pos = random_clustered(n, box, seed); mass = (1.0 + 0.25 * (np.arange(n) % 3)).astype(np.float32)
    P = sq.Params.defaults(box, nmesh, (box / n ** (1 / 3.)) / 30., TreeUseBH=1, TreeRcut=6.0)
    return box, nmesh, pos, mass, P


def _single(pos, mass, P):
    import shenqi_b200 as sq
    ctx = sq.Context(0)
    ctx.set_params(P)
    ctx.upload_particles(pos, mass); ctx.ph_sort(); ctx.tree_build()
    acc, pot, st = ctx.grav_short_tree()
    gpm, ppot = ctx.gravpm_force()
    keys, perm = ctx.sorted_keys_perm()
    ctx.close()
    return acc, pot, st.interactions, gpm, ppot, perm


def test_world1_equals_single_context():
    from shenqi_b200.dist import DistGravity, CudaBackend
    box, nmesh, pos, mass, P = _problem()
    acc, pot, nint, gpm, ppot, _ = _single(pos, mass, P)
    B = CudaBackend(P, 0)
    D = DistGravity(B)
    dpos, dmass = torch.from_numpy(pos).cuda(), torch.from_numpy(mass).cuda()
    a, p, st = D.tree_force(dpos, dmass)
    assert st.interactions == nint
    assert np.array_equal(a.cpu().numpy(), acc)            # same kernels, same order
    assert np.array_equal(p.cpu().numpy(), pot)
    out4 = D.pm_force(dpos, dmass).cpu().numpy()
    assert np.abs(out4[:, 1:] - gpm).max() <= 1e-9 * np.abs(gpm).max()
    assert np.abs(out4[:, 0] - ppot).max() <= 1e-9 * np.abs(ppot).max()
    B.close()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from shenqi_b200.dist import DistGravity, CudaBackend, Comm
        box, nmesh, pos, mass, P = _problem()
        acc, pot, nint, gpm, ppot, perm = _single(pos, mass, P) if rank == 0 else (None,) * 6
        import shenqi_b200 as sq
        c0 = sq.Context(rank); c0.set_params(P); c0.upload_particles(pos, mass); c0.ph_sort()
        _, order = c0.sorted_keys_perm(); c0.close()
        n = len(pos)
        cut = np.linspace(0, n, world + 1).astype(int); cut[1] = int(0.45 * n) if world == 2 else cut[1]
        mine = order[cut[rank]:cut[rank + 1]].astype(np.int64)
        np.random.default_rng(rank).shuffle(mine)
        B = CudaBackend(P, rank)
        D = DistGravity(B, Comm())
        dpos, dmass = torch.from_numpy(np.ascontiguousarray(pos[mine])).cuda(), torch.from_numpy(np.ascontiguousarray(mass[mine])).cuda()
        a, p, st = D.tree_force(dpos, dmass)
        out4 = D.pm_force(dpos, dmass)
        tot = torch.tensor([st.interactions], device="cuda"); dist.all_reduce(tot)
        res = {"a": a.cpu().numpy(), "p": p.cpu().numpy(), "out4": out4.cpu().numpy(), "mine": mine, "tot": int(tot.item())}
        # the overlapped step (PM chain on a second stream and communicator) must return the same numbers
        D2 = DistGravity(B, Comm(), comm_pm=Comm(dist.new_group()))
        for _ in range(2):
            a2, p2, o2, st2 = D2.step(dpos, dmass)
        torch.cuda.synchronize()
        res.update(step_same_tree=bool(torch.equal(a2, a) and torch.equal(p2, p) and st2.interactions == st.interactions),
                   step_pm_err=float((o2 - out4).abs().max() / out4.abs().max()))
        if rank == 0:
            res.update(acc=acc, pot=pot, nint=nint, gpm=gpm, ppot=ppot)
        q.put((rank, res))
        B.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_world2_nccl_equals_single_gpu():
    import torch.multiprocessing as mp
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = {}
    for _ in range(world):
        r, res = q.get(timeout=900)
        out[r] = res
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    ref = out[0]
    assert ref["tot"] == ref["nint"]
    for r in range(world):
        assert out[r]["step_same_tree"] and out[r]["step_pm_err"] <= 1e-12, (r, out[r]["step_same_tree"], out[r]["step_pm_err"])
    for r in range(world):
        m = out[r]["mine"]
        # every opening decision is the single-GPU one (interaction totals equal); the FP32 pair sums differ by rounding
        # because the groups of 32 start at other particles (other group centres)
        e = np.linalg.norm(out[r]["a"] - ref["acc"][m], axis=1) / np.linalg.norm(ref["acc"][m], axis=1)
        assert np.median(e) <= 1e-6 and e.max() <= 1e-4, (np.median(e), e.max())
        ep = np.abs(out[r]["p"] - ref["pot"][m]) / np.abs(ref["pot"][m])
        assert np.median(ep) <= 1e-6 and ep.max() <= 1e-4, (np.median(ep), ep.max())
        assert np.abs(out[r]["out4"][:, 1:] - ref["gpm"][m]).max() <= 1e-9 * np.abs(ref["gpm"]).max()
        assert np.abs(out[r]["out4"][:, 0] - ref["ppot"][m]).max() <= 1e-9 * np.abs(ref["ppot"]).max()
