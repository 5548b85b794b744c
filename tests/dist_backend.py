"""TEST INFRASTRUCTURE: a CPU stand-in for shenqi_b200.dist.CudaBackend built on the oracle (numpy + the C restatement),
so that the host logic of the multi-GPU step (routing, slab ownership, ghost planes, transposes, result return) can be
run on world_size-2 gloo without a GPU.  Never imported by the product."""
import ctypes
import numpy as np
import torch
from helpers import (oracle, ptr, o_keys, o_sort, o_tree, o_walk, o_cic_paint, o_cic_readout, c_int, c_dbl)


class OracleBackend:
    device = torch.device("cpu")

    def __init__(self, par):
        """par: dict of the oracle's GravParams (BoxSize, Nmesh, Asmth, G, ...)"""
        self.par = dict(par)
        self.box, self.nmesh = par["BoxSize"], par["Nmesh"]

    # ---- tree side
    def sort_local(self, pos, mass):
        p, m = pos.numpy(), mass.numpy()
        keys = o_keys(p, self.box)
        perm = o_sort(keys)
        return {"keys": torch.from_numpy(keys[perm].view(np.int64).copy()), "pos": torch.from_numpy(np.ascontiguousarray(p[perm])),
                "mass": torch.from_numpy(np.ascontiguousarray(m[perm])), "perm": torch.from_numpy(perm.astype(np.int64)), "geo": 0}

    def install_global(self, keys, pos, mass, geo):
        self.gpos = np.ascontiguousarray(pos.numpy()); self.gmass = np.ascontiguousarray(mass.numpy())
        self.oldacc = np.zeros(len(self.gpos))
        self.n = len(self.gpos)

    def tree_build(self):
        self.nodes, self.first = o_tree(self.gpos, self.gmass, self.box)
        return len(self.nodes)

    def set_oldacc_range(self, first, oldacc_sorted):
        self.oldacc[first:first + len(oldacc_sorted)] = oldacc_sorted.numpy()

    def walk_range(self, first, count):
        act = np.arange(first, first + count, dtype=np.int32)
        acc, pot, cnt, tot = o_walk(self.nodes, self.first, self.gpos, self.gmass, self.oldacc, self.par, active=act, want_pot=True)
        return torch.from_numpy(acc[first:first + count].copy()), torch.from_numpy(pot[first:first + count].copy()), int(tot)

    # ---- PM side (x-slabs; this stand-in needs at least 2 ranks, see slab_paint)
    def _cells(self, x):
        return (np.floor(x / (self.box / self.nmesh)).astype(np.int64)) % self.nmesh

    def route(self, pos, mass, P):
        p, m = pos.numpy(), mass.numpy()
        owner = self._cells(p[:, 0]) // (self.nmesh // P)
        order = np.argsort(owner, kind="stable")
        send4 = np.concatenate([p[order], m[order, None].astype(np.float64)], 1)
        slot = np.empty(len(p), np.int64); slot[order] = np.arange(len(p))
        return torch.from_numpy(send4), torch.from_numpy(slot), [int((owner == r).sum()) for r in range(P)]

    def unroute(self, back4, slot):
        return back4[slot]

    def new_slab(self, P, f):
        return torch.zeros((self.nmesh // P + 1, self.nmesh, self.nmesh), dtype=torch.float64)

    def slab_paint(self, part4, rank, P):
        assert P >= 2
        N, nxl = self.nmesh, self.nmesh // P
        a = part4.numpy()
        full = o_cic_paint(np.ascontiguousarray(a[:, :3]), np.ascontiguousarray(a[:, 3].astype(np.float32)), self.box, N)
        x0 = rank * nxl
        # every routed particle has its cell plane inside the slab, so the only other plane it touches is x0 + nxl
        others = np.ones(N, bool); others[x0:x0 + nxl] = False; others[(x0 + nxl) % N] = False
        assert not full[others].any()
        return torch.from_numpy(np.concatenate([full[x0:x0 + nxl], full[(x0 + nxl) % N][None]], 0))

    def fft_fwd(self, slab, P):
        N, nxl, nyl = self.nmesh, self.nmesh // P, self.nmesh // P
        a = np.fft.rfft2(slab.numpy()[:nxl], axes=(1, 2))
        return torch.from_numpy(np.ascontiguousarray(np.stack([a[:, q * nyl:(q + 1) * nyl] for q in range(P)], 0)))

    def fft_x(self, data, P, inverse):
        a = data.numpy()
        a[...] = np.fft.ifft(a, axis=0) * a.shape[0] if inverse else np.fft.fft(a, axis=0)

    def transfer(self, rho, rank, P):
        N, nyl, nzh = self.nmesh, self.nmesh // P, self.nmesh // 2 + 1
        full = np.zeros((N, N, nzh), np.complex128)
        full[:, rank * nyl:(rank + 1) * nyl] = rho.numpy()
        outs = [np.zeros_like(full) for _ in range(4)]
        oracle().sqo_pm_transfer(ptr(full), c_int(N), c_dbl(self.box), c_dbl(self.par["Asmth"]), c_dbl(self.par["G"]), *[ptr(o) for o in outs])
        return torch.from_numpy(np.ascontiguousarray(np.stack([o[:, rank * nyl:(rank + 1) * nyl] for o in outs], 0)))

    def transfer_phi(self, rho, rank, P):
        return self.transfer(rho, rank, P)[0].contiguous()

    def gradient(self, slab, lo2, hi3, P):
        """numpy statement of the 4-point difference on the slab extended by the neighbours' planes"""
        N, nxl = self.nmesh, self.nmesh // P
        phi = slab.numpy()
        ext = np.concatenate([lo2.numpy(), phi[:nxl], hi3.numpy()], 0)          # planes -2 .. nxl+2
        phi[nxl] = hi3.numpy()[0]
        scale = N / self.box
        def d(a, axis):
            if axis == 0:
                c = slice(2, 2 + nxl + 1)
                return -scale * (2. / 3 * (a[3:nxl + 4] - a[1:nxl + 2]) - 1. / 12 * (a[4:nxl + 5] - a[0:nxl + 1]))
            b = a[2:2 + nxl + 1]
            return -scale * (2. / 3 * (np.roll(b, -1, axis) - np.roll(b, 1, axis)) - 1. / 12 * (np.roll(b, -2, axis) - np.roll(b, 2, axis)))
        return [slab] + [torch.from_numpy(np.ascontiguousarray(d(ext, ax))) for ax in (0, 1, 2)]

    def fft_inv(self, recv, P, slab):
        N, nxl, nyl = self.nmesh, self.nmesh // P, self.nmesh // P
        r = recv.numpy()
        planes = np.concatenate([r[q] for q in range(P)], 1)          # [nxl, N, nzh]
        slab.numpy()[:nxl] = np.fft.irfft2(planes, s=(N, N), axes=(1, 2)) * float(N * N)

    def readout(self, part4, rank, P, slabs):
        N, nxl = self.nmesh, self.nmesh // P
        a = part4.numpy()
        pos = np.ascontiguousarray(a[:, :3])
        out = np.zeros((len(a), 4))
        x0 = rank * nxl
        for f, s in enumerate(slabs):
            full = np.zeros((N, N, N))
            full[x0:x0 + nxl] = s.numpy()[:nxl]
            full[(x0 + nxl) % N] = s.numpy()[nxl]
            out[:, f] = o_cic_readout(full, pos, self.box, N)
        return torch.from_numpy(out)
