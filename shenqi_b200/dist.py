"""Multi-GPU short-range TreePM step: one rank per GPU, exchanges over torch.distributed (NCCL on NVLink / NVSwitch;
gloo on CPU for the tests of this host logic).  SURVEY.md section 8e.

What is exchanged (and what the reference exchanges at the same place):
  * tree, production path (`DistGravity.tree_force_let`, everything behind the C ABI with raw NCCL, sq_let.cu): the tree is
    built under the reference's DomainDecomp - forced top-level nodes, own sub-trees, pseudo nodes - top-leaf moments are
    all-reduced (forcetree.cpp:1155-1282) and each rank imports, once per build, the nodes and particles of remote
    sub-trees that its particles can reach (instead of exporting boundary particles per walk, treewalk2.h:618-812);
    memory per GPU is O(N/P + boundary).
  * tree, replicated test mode (`tree_force`): every rank sorts its own particles, the sorted (key, pos, mass) ranges are all-gathered in rank order
    (= global Peano-Hilbert order, ranks own consecutive key ranges as after the reference's domain decomposition),
    every rank builds the whole tree and walks only its own targets.  The reference instead exchanges top-leaf
    moments (forcetree.cpp:1144-1282) and exports each boundary walk to the owner of the pseudo node and imports
    the partial result (treewalk2.h:282-373, 742-812); results are identical because the global tree is the same
    oct-tree.  The replicated tree costs O(N_total) memory and all-gather traffic per GPU (512^3 fits, 1024^3 does not).
  * PM: x-slabs.  Particles go to the owner of their x-plane (all-to-all of 32-byte records), are painted there,
    the one ghost plane goes to the next rank; the 3-D FFT is 2-D per plane, one all-to-all transpose, 1-D along x
    (forward once, backward for the potential and the three force components); results return by the reverse
    all-to-all.  The reference exchanges painted cells as z-pencils between domain owner and FFT owner
    (petapm.cpp:692-1107) around heFFTe's own transposes (petapm.cpp:281-303).

The compute between the exchanges is behind a small backend interface: `CudaBackend` (below) calls the C ABI of
libshenqi_b200.so with device pointers of torch tensors; the CPU tests pass a numpy backend built on the oracle.
"""
import torch
import torch.distributed as dist


class Comm:
    """The collectives the step needs, on whatever backend the default process group has."""

    def __init__(self, group=None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.nccl = dist.is_initialized() and dist.get_backend(group) == "nccl"
        self.bytes_sent = 0

    def allgather_ints(self, vals, device):
        """vals: list of python ints of this rank -> [world][len(vals)] list"""
        t = torch.tensor(vals, dtype=torch.int64, device=device)
        if self.world == 1:
            return [list(vals)]
        out = torch.empty(self.world * len(vals), dtype=torch.int64, device=device)
        dist.all_gather_into_tensor(out, t, group=self.group)
        return out.view(self.world, len(vals)).tolist()

    def allreduce_counts(self, counts, device):
        """sum of a list of python ints over the ranks (MPI_Allreduce of timebincounts, libgadget/timestep.cpp:378)"""
        if self.world == 1:
            return list(counts)
        t = torch.tensor(counts, dtype=torch.int64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t.tolist()

    def allgather_rows(self, t, counts):
        """t: [counts[rank], ...] -> concatenation over ranks in rank order"""
        if self.world == 1:
            return t
        mx = max(counts)
        pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
        out = torch.empty((self.world * mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, pad, group=self.group)
        self.bytes_sent += pad.numel() * pad.element_size() * (self.world - 1)
        if all(c == mx for c in counts):
            return out
        out = out.view((self.world, mx) + tuple(t.shape[1:]))
        return torch.cat([out[r, : counts[r]] for r in range(self.world)], 0)

    def alltoall_rows(self, send, send_counts, recv_counts):
        """send: rows grouped by destination rank (send_counts[r] rows to rank r) -> rows grouped by source rank"""
        if self.world == 1:
            return send
        recv = torch.empty((sum(recv_counts),) + tuple(send.shape[1:]), dtype=send.dtype, device=send.device)
        self.bytes_sent += (send.shape[0] - send_counts[self.rank]) * send[0:1].numel() * send.element_size()
        if self.nccl:
            dist.all_to_all_single(recv, send.contiguous(), list(recv_counts), list(send_counts), group=self.group)
            return recv
        self._p2p_alltoall(list(torch.split(send, list(send_counts))), list(torch.split(recv, list(recv_counts))))
        return recv

    def alltoall_equal(self, send):
        """send: [world, ...] block q goes to rank q -> [world, ...] block p came from rank p"""
        if self.world == 1:
            return send
        send = send.contiguous()
        recv = torch.empty_like(send)
        self.bytes_sent += send.numel() * send.element_size() * (self.world - 1) // self.world
        if self.nccl and not send.is_complex():
            dist.all_to_all_single(recv, send, group=self.group)
        elif self.nccl:
            dist.all_to_all_single(torch.view_as_real(recv), torch.view_as_real(send), group=self.group)
        else:
            self._p2p_alltoall([send[q] for q in range(self.world)], [recv[q] for q in range(self.world)])
        return recv

    def _p2p_alltoall(self, sends, recvs):
        # gloo has no all-to-all: pairwise exchange, lower rank sends first
        recvs[self.rank].copy_(sends[self.rank])
        for k in range(1, self.world):
            dst = (self.rank + k) % self.world
            src = (self.rank - k) % self.world
            s = sends[dst].contiguous()
            r = torch.empty_like(recvs[src]) if not recvs[src].is_contiguous() else recvs[src]
            if s.is_complex():
                s = torch.view_as_real(s); rr = torch.view_as_real(r)
            else:
                rr = r
            ops = []
            if s.numel():
                ops.append(dist.P2POp(dist.isend, s, dst, self.group))
            if rr.numel():
                ops.append(dist.P2POp(dist.irecv, rr, src, self.group))
            for w in (dist.batch_isend_irecv(ops) if ops else []):
                w.wait()
            if r is not recvs[src]:
                recvs[src].copy_(r)

    def shift(self, t, step):
        """send t to rank + step, return what rank - step sent (periodic)"""
        if self.world == 1:
            return t.clone()
        t = t.contiguous()
        out = torch.empty_like(t)
        dst, src = (self.rank + step) % self.world, (self.rank - step) % self.world
        self.bytes_sent += t.numel() * t.element_size()
        for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, t, dst, self.group), dist.P2POp(dist.irecv, out, src, self.group)]):
            w.wait()
        return out

    def barrier(self):
        if self.world > 1:
            dist.barrier(group=self.group)


class DistGravity:
    """One rank of the distributed force step.  `backend` does the local compute (CudaBackend or the tests' CPU one)."""

    def __init__(self, backend, comm=None, profile=False, comm_pm=None):
        self.B = backend
        self.comm = comm or Comm()
        # a second communicator lets step() run the PM exchanges concurrently with the tree's (NCCL serialises the
        # collectives of one communicator)
        self.comm_pm = comm_pm
        self.rank, self.world = self.comm.rank, self.comm.world
        self.profile = profile and torch.cuda.is_available()
        self._marks = []

    # ---- optional stage timing (CUDA events on the current stream; read with stage_ms())
    def _mark(self, name):
        if self.profile:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._marks.append((name, e))

    def stage_ms(self):
        """milliseconds between consecutive marks since the last call, summed by name"""
        torch.cuda.synchronize()
        out = {}
        for (n0, e0), (n1, e1) in zip(self._marks[:-1], self._marks[1:]):
            if n1 != "begin":
                out[n1] = out.get(n1, 0.0) + e0.elapsed_time(e1)
        self._marks = []
        return out

    # ------------------------------------------------------------------ short-range tree force
    def tree_force(self, pos, mass, oldacc=None, after_sort=None, before_walk=None):
        """pos [n,3] f64, mass [n] f32: this rank's particles (it must own a consecutive Peano-Hilbert key range, ranks in
        key order); oldacc [n] = |FullTreeGravAccel + GravPM| / G for the relative criterion or None.
        Returns (accel [n,3] already x G, potential [n] or None, stats) in the caller's particle order."""
        B, C = self.B, self.comm
        n = pos.shape[0]
        self._mark("begin")
        srt = B.sort_local(pos, mass)
        self._mark("tree:local_sort")
        self._sorted = srt          # PH-sorted copies of this rank's particles: pm_force(reuse_sorted=True) paints from them
        if after_sort is not None:
            after_sort()
        kmin = int(srt["keys"][0].item()) if n else 2 ** 63 - 1
        kmax = int(srt["keys"][-1].item()) if n else -1
        info = C.allgather_ints([n, kmin, kmax, int(srt["geo"])], B.device)
        counts = [r[0] for r in info]
        last = -1
        for r in info:      # keys are 63-bit, so the int64 view orders like the uint64 keys
            if r[0] and r[1] < last:
                raise ValueError("ranks must own consecutive Peano-Hilbert key ranges in rank order (domain decomposition)")
            if r[0]:
                last = r[2]
        self._mark("tree:counts_sync")
        gk = C.allgather_rows(srt["keys"], counts)
        gp = C.allgather_rows(srt["pos"], counts)
        gm = C.allgather_rows(srt["mass"], counts)
        self._mark("tree:allgather")
        B.install_global(gk, gp, gm, max(r[3] for r in info))
        self.tree_info = B.tree_build()
        self._mark("tree:global_build")
        first = sum(counts[: self.rank])
        perm = srt["perm"]
        if oldacc is not None:
            B.set_oldacc_range(first, oldacc[perm].contiguous())
        if before_walk is not None:
            before_walk()
        acc_s, pot_s, st = B.walk_range(first, n)
        self._mark("tree:walk")
        accel = torch.empty_like(acc_s)
        accel[perm] = acc_s
        pot = None
        if pot_s is not None:
            pot = torch.empty_like(pot_s)
            pot[perm] = pot_s
        self._mark("tree:unpermute")
        return accel, pot, st

    # ------------------------------------------------------------------ short-range tree force, domain-decomposed tree
    def tree_force_let(self, pos, mass, oldacc=None, after_sort=None, before_walk=None):
        """Like tree_force, but with the reference's partitioning: this rank's particles must be the ones the installed
        domain (backend.set_domain) assigns to it.  Local sort, tree under the domain's top nodes, moment exchange and
        sub-tree import (sq_tree_build_dist, NCCL inside the library), walk of the local particles.  No torch collective
        and no host list round trip on this path."""
        B = self.B
        n = pos.shape[0]
        self._mark("begin")
        T = B.tree
        T.upload_particles_device(pos.contiguous().data_ptr(), mass.contiguous().data_ptr(), n)
        if oldacc is not None:
            T.set_oldacc_device(oldacc.contiguous().data_ptr())
        T.ph_sort()
        self._mark("tree:local_sort")
        self._sorted = B.sorted_views(T, n)
        if after_sort is not None:
            after_sort()
        self.tree_info, self.let_info = T.tree_build_dist()
        self._mark("tree:build+moments+import")
        if before_walk is not None:
            before_walk()
        st = T.grav_short_tree_device()
        self._mark("tree:walk")
        accel = torch.empty((n, 3), dtype=torch.float64, device=B.device)
        pot = torch.empty(n, dtype=torch.float64, device=B.device)
        T.download_accel_device(accel.data_ptr(), pot.data_ptr())
        self._mark("tree:unpermute")
        return accel, pot, st

    # ------------------------------------------------------------------ particle-mesh force
    def pm_force(self, pos, mass, reuse_sorted=False, comm=None):
        """Returns out4 [n,4] = (PM potential, fx, fy, fz) for this rank's particles, in their order.
        reuse_sorted: pos/mass are the arrays the last tree_force() was called with; the PH-sorted copies it left are
        routed instead (PH-adjacent particles share mesh cells, so paint and readout stay cache-coherent, as on the
        single-GPU path and as after the reference's slots_gc_sorted) and the result is put back in the caller's order."""
        B, C = self.B, (comm or self.comm)
        perm = None
        if reuse_sorted and getattr(self, "_sorted", None) is not None and self._sorted["pos"].shape[0] == pos.shape[0]:
            pos, mass, perm = self._sorted["pos"], self._sorted["mass"], self._sorted["perm"]
        P, r = self.world, self.rank
        N = B.nmesh
        if N % P or N % 2:
            raise ValueError("Nmesh %d must be even and divisible by the number of ranks %d" % (N, P))
        nxl, nyl, nzh = N // P, N // P, N // 2 + 1
        self._mark("begin")
        send4, slot, counts = B.route(pos, mass, P)
        rc = C.allgather_ints(list(counts), B.device)
        recv_counts = [rc[q][r] for q in range(P)]
        self._mark("pm:route")
        part4 = C.alltoall_rows(send4, counts, recv_counts)
        self._mark("pm:particles_a2a")
        slab = B.slab_paint(part4, r, P)                       # [(nxl + 1), N, N]; last plane = ghost
        slab[0] += C.shift(slab[nxl], +1)                      # my ghost belongs to the next rank's first plane
        self._mark("pm:paint+ghost")
        packed = B.fft_fwd(slab, P)                            # [P, nxl, nyl, nzh] complex
        self._mark("pm:fft")
        rho = C.alltoall_equal(packed).view(N, nyl, nzh)       # block p holds planes p*nxl ..: x is the slowest index
        self._mark("pm:transpose_a2a")
        B.fft_x(rho, P, False)
        self._mark("pm:fft")
        phik = B.transfer_phi(rho, r, P)                       # [N, nyl, nzh]: phi_k
        self._mark("pm:transfer")
        B.fft_x(phik, P, True)
        self._mark("pm:fft")
        recv = C.alltoall_equal(phik.view(P, nxl, nyl, nzh))
        self._mark("pm:transpose_a2a")
        B.fft_inv(recv, P, slab)                               # the potential on my nxl planes
        self._mark("pm:fft")
        # forces = 4-point difference of the potential in real space (the reference's k-space filter, gravpm.cpp:447-475):
        # planes -2, -1 come from the previous rank, planes nxl .. nxl+2 (CIC ghost + stencil) from the next one
        lo2 = C.shift(slab[nxl - 2:nxl], +1)
        hi3 = C.shift(slab[0:3], -1)
        self._mark("pm:paint+ghost")
        slabs = B.gradient(slab, lo2, hi3, P)                  # [phi, fx, fy, fz], each (nxl + 1) planes
        self._mark("pm:gradient")
        res = B.readout(part4, r, P, slabs)
        self._mark("pm:readout")
        back = C.alltoall_rows(res, recv_counts, counts)
        self._mark("pm:particles_a2a")
        out = B.unroute(back, slot)
        if perm is not None:
            res = torch.empty_like(out)
            res[perm] = out
            out = res
        self._mark("pm:route")
        return out


    # ------------------------------------------------------------------ one PM step: both forces, chains overlapped
    def step(self, pos, mass, oldacc=None, pm_host_out=None):
        """tree_force + pm_force of one PM step (run.cpp:538-563).  With a backend that has a second stream and a second
        communicator the PM chain (route, all-to-all, paint, FFTs, transposes, readout) runs on that stream from the
        moment the local sort is done, underneath the tree chain (all-gather, global build); both join before the
        walk, which wants every SM.  pm_host_out: pinned host tensor [n, 4]; the PM result is also copied there on the
        second stream, underneath the walk.  Returns (accel, potential, out4, stats) like the two calls."""
        B = self.B
        side = getattr(B, "pm_stream", None)
        tree_force = self.tree_force_let if getattr(B, "tree", None) is not None else self.tree_force
        if side is None or self.comm_pm is None or self.profile or self.world == 1:
            a, p, st = tree_force(pos, mass, oldacc)
            o = self.pm_force(pos, mass, reuse_sorted=True)
            if pm_host_out is not None:
                pm_host_out.copy_(o, non_blocking=True)
            return a, p, o, st
        main = torch.cuda.current_stream(B.device)
        box = {}

        def fork():
            ev = torch.cuda.Event()
            ev.record(main)
            with torch.cuda.stream(side):
                side.wait_event(ev)
                box["out"] = self.pm_force(pos, mass, reuse_sorted=True, comm=self.comm_pm)
                box["done"] = torch.cuda.Event()
                box["done"].record(side)
                if pm_host_out is not None:
                    pm_host_out.copy_(box["out"], non_blocking=True)      # a DMA: runs while the walk owns the SMs

        def join():
            main.wait_event(box["done"])
            box["out"].record_stream(main)

        a, p, st = tree_force(pos, mass, oldacc, after_sort=fork, before_walk=join)
        return a, p, box["out"], st


class CudaBackend:
    """Local compute of DistGravity on one B200 through the C ABI (device pointers of torch tensors).
    Two library contexts: `loc` sorts and routes this rank's own particles, `glob` holds the all-gathered set, the tree
    and the walk.  Both run on torch's current stream, so the exchanges and the kernels are ordered without host syncs."""

    def __init__(self, params, device, nccl_id=None, rank=0, nranks=1):
        """nccl_id (comm_unique_id() of rank 0, the same 128 bytes on every rank): also create `tree`, the context of the
        domain-decomposed tree path, as rank `rank` of `nranks` (collective)."""
        from . import api
        self.api = api
        self.device = torch.device("cuda", device) if isinstance(device, int) else device
        torch.cuda.set_device(self.device)
        self.loc = api.Context(self.device.index)
        self.glob = api.Context(self.device.index)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        self.loc.set_stream(stream); self.glob.set_stream(stream)
        self.tree = None
        if nccl_id is not None:
            self.tree = api.Context(self.device.index, nccl_id=nccl_id, rank=rank, nranks=nranks)
            self.tree.set_stream(stream)
        # PM kernels have their own context; DistGravity.step() moves it to `pm_stream` while the tree chain runs
        self.pmc = api.Context(self.device.index)
        self.pmc.set_stream(stream)
        self.pm_stream = torch.cuda.Stream(self.device)
        self.set_params(params)
        self._slabs = {}

    def set_params(self, params):
        self.params = params
        self.nmesh = params.Nmesh
        self.loc.set_params(params)
        self.glob.set_params(params)
        self.pmc.set_params(params)
        if self.tree is not None:
            self.tree.set_params(params)

    def set_domain(self, dom):
        self.tree.domain_set(dom)

    def sorted_views(self, ctx, n):
        A = self.api
        # copies: the library may move its particle arrays when the imported particles are appended behind them
        return {"pos": self._wrap(ctx.device_ptr(A.PTR_SORTED_POS)[0], (n, 3), torch.float64).clone(),
                "mass": self._wrap(ctx.device_ptr(A.PTR_SORTED_MASS)[0], (n,), torch.float32).clone(),
                "perm": self._wrap(ctx.device_ptr(A.PTR_SORTED_PERM)[0], (n,), torch.int32).to(torch.int64)}

    def close(self):
        self.loc.close(); self.glob.close(); self.pmc.close()
        if self.tree is not None:
            self.tree.close()

    def _wrap(self, ptr, shape, dtype):
        """torch view of library-owned device memory (valid until the library reallocates it)"""
        n = 1
        for s in shape:
            n *= s
        esz = torch.empty(0, dtype=dtype).element_size()

        class _Mem:      # __cuda_array_interface__ holder
            pass
        m = _Mem()
        typestr = {torch.float64: "<f8", torch.float32: "<f4", torch.int64: "<i8", torch.int32: "<i4"}[dtype]
        m.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False), "version": 2, "strides": None}
        return torch.as_tensor(m, device=self.device) if n else torch.empty(shape, dtype=dtype, device=self.device)

    # ---- tree side
    def sort_local(self, pos, mass):
        """this rank's particles (host or device tensors) -> PH-sorted device tensors (views of library memory)"""
        n = pos.shape[0]
        if pos.is_cuda:
            self.loc.upload_particles_device(pos.contiguous().data_ptr(), mass.contiguous().data_ptr(), n)
        else:
            self.loc.upload_particles(pos.numpy(), mass.numpy())
        self.loc.ph_sort()
        A = self.api
        keys = self._wrap(self.loc.device_ptr(A.PTR_SORTED_KEYS)[0], (n,), torch.int64)
        spos = self._wrap(self.loc.device_ptr(A.PTR_SORTED_POS)[0], (n, 3), torch.float64)
        smass = self._wrap(self.loc.device_ptr(A.PTR_SORTED_MASS)[0], (n,), torch.float32)
        perm = self._wrap(self.loc.device_ptr(A.PTR_SORTED_PERM)[0], (n,), torch.int32).to(torch.int64)
        # "geometric path differs from the key digits" (particles exactly on cell faces): carried to the global tree
        return {"keys": keys, "pos": spos, "mass": smass, "perm": perm, "geo": self.loc.geo_mismatch()}

    def install_global(self, keys, pos, mass, geo):
        self.glob.set_sorted_device(keys.data_ptr(), pos.data_ptr(), mass.data_ptr(), keys.shape[0], geo)

    def tree_build(self):
        return self.glob.tree_build()

    def set_oldacc_range(self, first, oldacc_sorted):
        self.glob.set_oldacc_range_device(first, oldacc_sorted.shape[0], oldacc_sorted.data_ptr())

    def walk_range(self, first, count):
        st = self.glob.tree_walk_range(first, count)
        A = self.api
        n = self.glob.n
        acc = self._wrap(self.glob.device_ptr(A.PTR_WALK_ACCEL)[0], (n, 3), torch.float64)[first:first + count]
        try:
            pot = self._wrap(self.glob.device_ptr(A.PTR_WALK_POT)[0], (n,), torch.float64)[first:first + count]
        except A.ShenqiError:
            pot = None
        return acc, pot, st

    # ---- PM side
    def _pm(self):
        """the PM context, on whatever stream is current (cuFFT plans follow the stream)"""
        st = torch.cuda.current_stream(self.device).cuda_stream
        if st != getattr(self, "_pm_bound", None):
            self.pmc.set_stream(st)
            self._pm_bound = st
        return self.pmc

    def route(self, pos, mass, P):
        n = pos.shape[0]
        send4 = torch.empty((n, 4), dtype=torch.float64, device=self.device)
        slot = torch.empty(n, dtype=torch.int32, device=self.device)
        counts = self._pm().pm_slab_route(pos.data_ptr(), mass.data_ptr(), n, P, send4.data_ptr(), slot.data_ptr())
        return send4, slot, counts

    def unroute(self, back4, slot):
        out = torch.empty_like(back4)
        self._pm().pm_slab_unroute(back4.data_ptr(), slot.data_ptr(), slot.shape[0], out.data_ptr())
        return out

    def new_slab(self, P, f):
        N = self.nmesh
        return self._buf("slab%d" % f, (N // P + 1, N, N), torch.float64)

    def _buf(self, name, shape, dtype):
        t = self._slabs.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = torch.empty(shape, dtype=dtype, device=self.device)
            self._slabs[name] = t
        return t

    def slab_paint(self, part4, rank, P):
        N = self.nmesh
        slab = self._buf("slab0", (N // P + 1, N, N), torch.float64)
        self._pm().pm_slab_paint(part4.data_ptr(), part4.shape[0], rank, P, slab.data_ptr())
        return slab

    def fft_fwd(self, slab, P):
        N = self.nmesh
        nxl, nzh = N // P, N // 2 + 1
        work = self._buf("work", (nxl, N, nzh), torch.complex128)
        packed = self._buf("packed", (P, nxl, N // P, nzh), torch.complex128)
        self._pm().pm_slab_fft_fwd(slab.data_ptr(), P, work.data_ptr(), packed.data_ptr())
        return packed

    def fft_x(self, data, P, inverse):
        assert data.is_contiguous()
        self._pm().pm_slab_fft_x(data.data_ptr(), P, inverse)

    def transfer(self, rho, rank, P):
        N = self.nmesh
        out4 = self._buf("out4", (4, N, N // P, N // 2 + 1), torch.complex128)
        self._pm().pm_slab_transfer(rho.data_ptr(), rank, P, out4.data_ptr())
        return out4

    def transfer_phi(self, rho, rank, P):
        N = self.nmesh
        phik = self._buf("phik", (N, N // P, N // 2 + 1), torch.complex128)
        self._pm().pm_slab_transfer_phi(rho.data_ptr(), rank, P, phik.data_ptr())
        return phik

    def gradient(self, slab, lo2, hi3, P):
        N = self.nmesh
        ext = self._buf("ext", (N // P + 5, N, N), torch.float64)
        f = [self.new_slab(P, k) for k in (1, 2, 3)]
        self._pm().pm_slab_gradient(slab.data_ptr(), lo2.contiguous().data_ptr(), hi3.contiguous().data_ptr(), P, ext.data_ptr(),
                                  f[0].data_ptr(), f[1].data_ptr(), f[2].data_ptr())
        return [slab] + f

    def fft_inv(self, recv, P, slab):
        N = self.nmesh
        work = self._buf("work", (N // P, N, N // 2 + 1), torch.complex128)
        self._pm().pm_slab_fft_inv(recv.data_ptr(), P, work.data_ptr(), slab.data_ptr())

    def readout(self, part4, rank, P, slabs):
        out = torch.empty_like(part4)
        self._pm().pm_slab_readout(part4.data_ptr(), part4.shape[0], rank, P, [s.data_ptr() for s in slabs], out.data_ptr())
        return out


def global_pushdown_bin(comm, local_counts, largest_active, isPM, device="cpu"):
    """libgadget/timestep.cpp:378-398 across ranks: all-reduce this rank's timebincounts (the output of
    Context.gravity_timebins_device), then the largest occupied bin and the PM-step push-down rule on the GLOBAL counts,
    so that every rank clamps to the same bin (Context.timebin_pushdown_device).  Returns (push_down_bin, global counts)."""
    from .api import pushdown_bin
    return pushdown_bin(comm.allreduce_counts(local_counts, device), largest_active, isPM)
