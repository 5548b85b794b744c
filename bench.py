#!/usr/bin/env python
"""Headline benchmark (BASELINE.json): short-range TreePM gravity, particles/s and interactions/s.

One "step" = one full pass of the hot path over the particle set, device-resident:
   PH keys -> radix sort -> gather -> tree build -> moments -> group walk (relative criterion)
   -> PM: CIC paint -> r2c -> Green/differencing -> 4 x c2r -> fused CIC readout.
N=1 workload = BASELINE.json configs[1]: synthetic Zel'dovich 256^3 DM particles, Nmesh=512, z=9.
`value` = particles whose short-range + PM accelerations were produced per second (all ranks).
`e2e`   = the same through the C ABI with HOST buffers: H2D of positions/masses/old accelerations and D2H of
          the tree and PM accelerations inside the timed region (pinned host memory).
`--impl reference` times the reference's own CPU code (oracle/_ref, OpenMP, all host cores) on a bounded
sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import numpy as np

# 23 flop per interaction, 28 when the potential is accumulated as well.  The step benchmarked is a
# PM step on a full-particle tree, where the reference also updates the potential (gravshort-tree2.cpp:148).
FLOP_PER_INTERACTION = 28
# dram__bytes_read.sum + dram__bytes_write.sum of one sq_walk_kernel launch on the N=1 workload (256^3), from the
# `ncu --set full` capture summarised in profiles/r1_walk_v5c_sections.txt (1.663 GB + 0.526 GB); None for other sizes
WALK_DRAM_TRAFFIC_256 = 2.176e9     # ncu --set full, walk v6b at 256^3: dram__bytes_read.sum 1.650 GB + dram__bytes_write.sum 0.526 GB (profiles/r1_walk_v6_sections.txt)
DIST_NGRID = {2: 320, 4: 400, 8: 512}   # ~256^3 particles per GPU (weak scaling), Nmesh = 2 Ngrid divisible by the ranks


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def workload(ngrid, nmesh):
    from shenqi_b200.ics import zeldovich, rho0, G_INTERNAL
    box = 1000.0 * ngrid            # 1 Mpc/h mean spacing, kpc/h units
    pos, mass, ids = zeldovich(ngrid, box, rms_disp=0.2, seed=181170)
    par = dict(BoxSize=box, Nmesh=nmesh, GravitySoftening=(box / ngrid) / 30., G=G_INTERNAL, rho0=rho0(), TreeRcut=6.0,
               ErrTolForceAcc=0.002, BHOpeningAngle=0.175, MaxBHOpeningAngle=0.9)
    return pos, mass, par


def pinned(a):
    """copy of a numpy array in page-locked host memory (torch is only the allocator here)."""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t.numpy(), t


# ---------------------------------------------------------------------------------------------- CPU side
def cpu_threads():
    """all host cores for the reference's OpenMP regions, also under torchrun (which exports OMP_NUM_THREADS=1)."""
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        pass
    return n


def reference_pm_subprocess(pos, mass, box, nmesh, G, asmth=1.5, timeout=900):
    """tools/ref_pm_once.py on (pos, mass): {"pm_ref_s": seconds of gravpm_force, "gravpm_ref": accelerations in input order}
    or {"pm_ref_s": None, "pm_ref_error": why}."""
    import subprocess, tempfile
    scratch = os.path.join(ROOT, "oracle", "_ref", "scratch")
    try:
        os.makedirs(scratch, exist_ok=True)
        with tempfile.TemporaryDirectory(dir=scratch) as d:
            src, dst = os.path.join(d, "in.npz"), os.path.join(d, "out.npz")
            np.savez(src, pos=pos, mass=mass, box=np.float64(box), nmesh=np.int64(nmesh), asmth=np.float64(asmth), G=np.float64(G))
            env = dict(os.environ, OMP_NUM_THREADS=str(cpu_threads()))
            for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
                env.pop(k, None)
            r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ref_pm_once.py"), src, dst], env=env, timeout=timeout,
                               stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
            if r.returncode != 0 or not os.path.exists(dst):
                return {"pm_ref_s": None, "pm_ref_error": "exit %d: %s" % (r.returncode, r.stderr.decode(errors="replace")[-300:])}
            o = np.load(dst)
            return {"pm_ref_s": float(o["seconds"]), "gravpm_ref": o["gravpm"]}
    except Exception as e:                                   # noqa
        return {"pm_ref_s": None, "pm_ref_error": repr(e)}


def cpu_reference_step(pos, mass, par, old=None, want="auto", with_pm=True):
    """One step of the reference's own CPU code (oracle/_ref: slots_gc_sorted, force_tree_full, grav_short_tree, OpenMP, all
    host cores) or, if that was not built, of the C port (oracle/).  old = (FullTreeGravAccel, GravPM) inputs of the
    relative criterion in input order (None: bootstrapped by an untimed Barnes-Hut pass).  The PM is timed separately and is
    NOT part of step_s: pm_s = the numpy restatement, pm_ref_s = the reference's own gravpm_force (oracle/_ref compiles
    petapm.cpp against a stand-in for the absent FFTW/heFFTe, so its FFT time is not the reference's).
    Returns dict of seconds, interaction count and the accelerations in input order."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as H
    n = len(pos)
    box = par["BoxSize"]
    meansep = box / round(n ** (1 / 3.))
    opar = H.default_params(box, par["Nmesh"], n, G=par["G"], TreeUseBH=0, TreeRcut=par["TreeRcut"], GravitySoftening=par["GravitySoftening"], rho0=par["rho0"])
    out = {}
    use_ref = H.ref() is not None and want != "port"
    if use_ref:
        H.ref().ref_set_threads(cpu_threads())
    with H.quiet_stdout():
        return _cpu_reference_step(H, pos, mass, par, opar, meansep, use_ref, out, old, with_pm)


def _cpu_reference_step(H, pos, mass, par, opar, meansep, use_ref, out, old, with_pm):
    n = len(pos)
    box = par["BoxSize"]
    if use_ref:
        R = H.ref()
        ids = np.arange(1, n + 1, dtype=np.uint64)
        H.r_load(pos, mass, box, ids=ids)
        t0 = time.perf_counter(); R.ref_sort_particles(); out["sort_s"] = time.perf_counter() - t0
        _, _, rids = H.r_sorted_particles()
        inv = (rids - 1).astype(np.int64)
        t0 = time.perf_counter(); R.ref_tree_build(H.c_dbl(0.9), None); out["tree_s"] = time.perf_counter() - t0
        if old is None:
            # bootstrap |a_old| with one Barnes-Hut pass (not timed), then the timed relative-criterion walk
            H.r_walk(dict(opar, TreeUseBH=1), meansep)
            fa = ga = None
        else:
            fa = np.ascontiguousarray(old[0][inv]); ga = np.ascontiguousarray(old[1][inv])
            R.ref_set_accels(H.ptr(fa), H.ptr(ga))
        _, tot = H.r_counts(opar, meansep)
        t0 = time.perf_counter(); acc, _ = H.r_walk(opar, meansep, fulltree=fa, gravpm=ga); out["walk_s"] = time.perf_counter() - t0
        out["interactions"] = int(tot)
        out["kind"] = "reference"
        out["cores"] = R.ref_omp_threads()
        a = np.zeros((n, 3)); a[inv] = acc
        out["accel"] = a
        R.ref_tree_free()
    else:
        t0 = time.perf_counter(); perm = H.o_sort(H.o_keys(pos, box)); spos = np.ascontiguousarray(pos[perm]); smass = np.ascontiguousarray(mass[perm]); out["sort_s"] = time.perf_counter() - t0
        t0 = time.perf_counter(); nodes, first = H.o_tree(spos, smass, box); out["tree_s"] = time.perf_counter() - t0
        if old is None:
            a1, _, _, _ = H.o_walk(nodes, first, spos, smass, np.zeros(n), dict(opar, TreeUseBH=1))
            oldacc = np.linalg.norm(a1, axis=1) / par["G"]
        else:
            oldacc = np.linalg.norm(old[0][perm] + old[1][perm], axis=1) / par["G"]
        t0 = time.perf_counter(); acc, _, _, tot = H.o_walk(nodes, first, spos, smass, oldacc, opar); out["walk_s"] = time.perf_counter() - t0
        out["interactions"] = int(tot)
        out["kind"] = "port"
        out["cores"] = os.cpu_count()
        a = np.zeros((n, 3)); a[perm] = acc
        out["accel"] = a
    out["pm_s"] = out["pm_ref_s"] = None
    if with_pm:
        t0 = time.perf_counter(); H.o_pm_force(pos, mass, box, par["Nmesh"], 1.5, par["G"]); out["pm_s"] = time.perf_counter() - t0
        if use_ref and hasattr(H.ref(), "ref_petapm_fft"):
            # the reference's own gravpm_force (gravpm.cpp + petapm.cpp, unmodified; includes the tree it builds for its
            # regions; its FFT is the stand-in of oracle/ref_shim/heffte.h, not FFTW) - in a process of its own: an
            # abort inside the reference must not take the benchmark down.  A reported extra, never fatal.
            out.update(reference_pm_subprocess(pos, mass, box, par["Nmesh"], par["G"]))
    out["step_s"] = out["sort_s"] + out["tree_s"] + out["walk_s"]
    return out


def bench_config(ngrid, nmesh, n, group_width):
    """`config` of the JSON line: identical for both arms (the reference arm runs on OUR arm's config)."""
    return {"workload": "zeldovich-%d^3-nmesh%d-z9" % (ngrid, nmesh), "particles": n, "Nmesh": nmesh, "ErrTolForceAcc": 0.002, "TreeRcut": 6.0,
            "Asmth": 1.5, "group_width": group_width,
            "l2": "inputs_exceed_l2 (particle arrays %.0f MB, mesh %.0f MB)" % (n * 36 / 1e6, float(nmesh) ** 3 * 8 / 1e6)}


def run_reference_arm(args):
    """`--impl reference`: the reference's own CPU implementation of the path (oracle/_ref = its unmodified sources:
    slots_gc_sorted + force_tree_full + grav_short_tree, OpenMP over all host cores) on OUR arm's config.  Each step is a
    bounded sample of that workload: a periodic cpu_ngrid^3 sub-volume with the same mean spacing, softening, Rcut,
    Nmesh/Ngrid ratio and ErrTolForceAcc (the full 256^3 step takes ~40 s on 16 cores; `--cpu-ngrid 256` runs it).  value =
    particles per second of sort + tree + walk, reference code only; the PM restatement (numpy) is timed once and reported
    under cpu_baseline.pm_restatement_s, outside value.  Under torchrun rank 0 alone runs, with every host core."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    os.environ["OMP_NUM_THREADS"] = str(cpu_threads())
    ngrid = args.cpu_ngrid
    nmesh = args.nmesh or 2 * args.ngrid
    pos, mass, par = workload(ngrid, ngrid * nmesh // args.ngrid)
    n = len(pos)
    times = []
    res = None
    for it in range(args.warmup + args.steps):
        res = cpu_reference_step(pos, mass, par, with_pm=False)
        if it >= args.warmup:
            times.append(res["step_s"])
    pmres = cpu_reference_step(pos, mass, par, with_pm=True) if ngrid <= 128 else {"pm_s": None, "pm_ref_s": None}
    pm = pmres["pm_s"]
    t = float(np.mean(times))
    v = n / t
    sample = ("each step = a periodic %d^3 / Nmesh %d sub-volume of the workload (1/%d of its particles; same spacing, softening, Rcut, ErrTolForceAcc): "
              "PH sort + tree build + relative-criterion walk by the %s; particles/s of that" % (ngrid, ngrid * nmesh // args.ngrid, (args.ngrid // ngrid) ** 3,
              "reference's own OpenMP code (oracle/_ref)" if res["kind"] == "reference" else "C port (oracle/)"))
    line = {"impl": "reference", "metric": "short-range gravity particles/s", "value": v, "unit": "particles/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": bench_config(args.ngrid, nmesh, args.ngrid ** 3, args.group_width),
            "interactions_per_s": res["interactions"] / res["walk_s"],
            "cpu_baseline": {"value": v, "unit": "particles/s", "cores": res["cores"], "kind": res["kind"], "sample": sample,
                             "seconds": {k: res[k] for k in ("sort_s", "tree_s", "walk_s")}, "pm_restatement_s (numpy, not in value)": pm,
                             "pm_reference_s (the reference's gravpm_force with a stand-in FFT instead of FFTW: an upper bound, not in value)": pmres["pm_ref_s"]},
            "e2e": {"value": v, "unit": "particles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------- multi-GPU
def torch_topleaf(dom, keys):
    """domain_get_topleaf (domain.h:68-75) for a device tensor of keys (int64 view of the 63-bit keys)."""
    import torch
    dev = keys.device
    start = torch.from_numpy(dom.topnodes["StartKey"].astype(np.int64)).to(dev)
    daughter = torch.from_numpy(dom.topnodes["Daughter"].astype(np.int64)).to(dev)
    shift = torch.from_numpy(dom.topnodes["Shift"].astype(np.int64)).to(dev)
    leaf = torch.from_numpy(dom.topnodes["Leaf"].astype(np.int64)).to(dev)
    no = torch.zeros_like(keys)
    for _ in range(22):
        d = daughter[no]
        live = d >= 0
        if not bool(live.any()):
            break
        sub = (keys - start[no]) >> (shift[no] - 3).clamp(min=0)
        no = torch.where(live, d + sub, no)
    return leaf[no]


def run_dist(args, rank, local_rank, world):
    """N > 1: ONE domain-decomposed problem (not replicas): ~256^3 particles per GPU of a global box.  Partitioning = the
    reference's: a top tree over Peano-Hilbert key ranges whose leaves are cut into one contiguous, equal-cost segment per
    rank (sq_domain_from_histogram; cost = particle count, or the measured interaction cost with --balance cost).
    Step = shenqi_b200.dist.DistGravity.step: local sort, tree under the domain's top nodes, top-leaf moment all-reduce,
    import of the reachable remote sub-trees (raw NCCL inside the library), walk of the local particles; PM on x-slabs with
    all-to-all routing / transposes on a second stream and communicator underneath the tree chain."""
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":              # NCCL prints its banner on stdout: keep it to the one JSON line
        del os.environ["NCCL_DEBUG"]
    # a rank that dies (e.g. out of memory) leaves the others waiting inside a collective: give up after --deadline seconds
    # instead of holding 8 GPUs until somebody's timeout (a 1024^3 attempt of round 2 did exactly that for 15 minutes)
    import signal
    import datetime
    signal.signal(signal.SIGALRM, signal.SIG_DFL)       # default action = the kernel ends the process, even inside a CUDA / NCCL call
    signal.alarm(max(60, int(args.deadline)))
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=max(60, int(args.deadline))))
    import shenqi_b200 as sq
    from shenqi_b200.ics import zeldovich_torch, clustered_torch, rho0, G_INTERNAL
    from shenqi_b200.dist import DistGravity, CudaBackend, Comm

    ngrid = args.ngrid if args.ngrid != 256 else DIST_NGRID.get(world, int(round(256 * world ** (1 / 3.) / 8)) * 8)
    nmesh = args.nmesh or 2 * ngrid
    W = max(args.warmup, 3)
    clustered = args.ics == "clustered"
    box = args.box if args.box > 0 else (50000.0 * ngrid / 512 if clustered else 1000.0 * ngrid)   # C3: 512^3 in 50 Mpc/h
    P = sq.Params.defaults(box, nmesh, (box / ngrid) / 30., G=G_INTERNAL, rho0=rho0(), TreeRcut=6.0, TreeUseBH=1)
    # the library's own communicator: ncclGetUniqueId on rank 0, handed to the others (MPI_Bcast in the reference's host)
    idt = torch.zeros(sq.api.NCCL_UNIQUE_ID_BYTES, dtype=torch.uint8, device=dev)
    if rank == 0:
        idt.copy_(torch.from_numpy(sq.comm_unique_id()))
    dist.broadcast(idt, 0)
    B = CudaBackend(P, local_rank, nccl_id=idt.cpu().numpy(), rank=rank, nranks=world)
    B.tree.set_group_width(args.group_width)
    D = DistGravity(B, Comm(), profile=False, comm_pm=Comm(dist.new_group()))
    fp32_peak = B.tree.measure_fp32_peak()
    hbm_peak, peak_src = read_peaks()

    def allmax(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allmin(x):
        return -allmax(-x)

    def allsum(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- set-up (untimed): global ICs on every GPU (a base realisation, periodically tiled T^3 times for the sizes that no
    # single GPU can generate: 1024^3 = 2^3 tiles of a 512^3 box); domain from the key histogram; every rank keeps the
    # particles the domain gives it (what the reference's domain exchange leaves on a rank)
    T = max(1, args.tile)
    nbase, bbox = ngrid // T, box / T
    if clustered:
        base_pos, base_mass, ics_tag = clustered_torch(nbase, bbox, dev, frac=args.clustered_frac)
    else:
        base_pos, base_mass = zeldovich_torch(nbase, bbox, dev)
        ics_tag = "zeldovich rms 0.2 spacings, seed 181170"
    if T > 1:
        base_pos.clamp_(min=bbox * 1e-9)      # no coordinate exactly on a tile boundary (= a cell face of the top levels)
        ics_tag += ", %d^3 base realisation tiled %d^3 times" % (nbase, T)
    ntot = base_pos.shape[0] * T ** 3
    level = args.domain_level

    def tiles():
        for t in range(T ** 3):
            off = torch.tensor([(t // (T * T)) * bbox, ((t // T) % T) * bbox, (t % T) * bbox], dtype=torch.float64, device=dev)
            pt = base_pos if T == 1 else (base_pos + off)
            kt = torch.empty(pt.shape[0], dtype=torch.int64, device=dev)
            B.tree.keys_device(pt.data_ptr(), pt.shape[0], kt.data_ptr())
            yield t, pt, kt

    def make_domain(cost_tiles=None):
        cnt = torch.zeros(8 ** level, dtype=torch.int64, device=dev)
        cst = None if cost_tiles is None else torch.zeros(8 ** level, dtype=torch.float64, device=dev)
        for t, pt, kt in tiles():
            cell = kt >> (63 - 3 * level)
            cnt += torch.bincount(cell, minlength=8 ** level)
            if cst is not None:
                cst += torch.bincount(cell, weights=cost_tiles[t], minlength=8 ** level)
        return sq.domain_from_histogram(cnt.cpu().numpy(), world, args.overdecomposition, cost=None if cst is None else cst.cpu().numpy())

    def take(dom):
        task = torch.from_numpy(dom.topleaves["Task"].astype(np.int64)).to(dev)
        sel, ps, ms = [], [], []
        for t, pt, kt in tiles():
            mine = torch.nonzero(task[torch_topleaf(dom, kt)] == rank).flatten()
            sel.append(mine + t * base_pos.shape[0]); ps.append(pt[mine]); ms.append(base_mass[mine])
        mine, pos, mass = torch.cat(sel), torch.cat(ps), torch.cat(ms)
        g = torch.Generator(device=dev); g.manual_seed(1234 + rank)
        sh = torch.randperm(mine.shape[0], device=dev, generator=g)
        return mine[sh], pos[sh].contiguous(), mass[sh].contiguous()

    def bootstrap(pos, mass):
        # |a_old| from a Barnes-Hut pass + PM, as the reference starts (gravshort-tree2.cpp:170-171)
        P.TreeUseBH = 1
        B.set_params(P)
        a0, _, st0 = D.tree_force_let(pos, mass)
        o4 = D.pm_force(pos, mass)
        oldacc = ((a0 + o4[:, 1:4]).norm(dim=1) / G_INTERNAL).contiguous()
        P.TreeUseBH = 0
        B.set_params(P)
        return a0, oldacc, st0

    dom = make_domain()
    B.set_domain(dom)
    mine, pos, mass = take(dom)
    a0, oldacc, st0 = bootstrap(pos, mass)
    balance = {"by": "particle count"}
    if args.balance == "cost":
        # the reference balances the measured work (GravCost, domain.cpp:452-527): one relative-criterion walk, its
        # per-particle interaction cost summed per histogram cell, new segments of equal cost, particles re-dealt
        D.tree_force_let(pos, mass, oldacc)
        cost_all = torch.zeros(ntot, dtype=torch.float32, device=dev)
        cmine = torch.empty(pos.shape[0], dtype=torch.float32, device=dev)
        B.tree.walk_particle_cost(cmine.data_ptr())
        cost_all[mine] = cmine
        dist.all_reduce(cost_all)
        nb = base_pos.shape[0]
        dom = make_domain([cost_all[t * nb:(t + 1) * nb].double() for t in range(T ** 3)])
        B.set_domain(dom)
        del cost_all, cmine
        mine, pos, mass = take(dom)
        a0, oldacc, st0 = bootstrap(pos, mass)
        balance = {"by": "interaction cost of the previous walk"}
    n = pos.shape[0]

    # ---- parity against ONE rank on the same domain (rank 0, untimed): Barnes-Hut pass, interaction totals equal, |da|/|a|
    verify = None
    if args.verify and T == 1 and ntot <= 200000000:
        pos_all, mass_all = base_pos, base_mass
        tot_bh = allsum(float(st0.interactions))
        if rank == 0:
            d1 = sq.Domain(dom.topnodes.copy(), dom.topleaves.copy(), np.array([(0, dom.ntopleaves)], sq.TASK_DTYPE))
            d1.topleaves["Task"] = 0
            torch.cuda.synchronize()
            c1 = sq.Context(local_rank)
            P.TreeUseBH = 1
            c1.set_params(P)
            c1.upload_particles_device(pos_all.data_ptr(), mass_all.data_ptr(), ntot)
            c1.ph_sort(); c1.domain_set(d1); c1.tree_build_dist()
            st1 = c1.grav_short_tree_device()
            a1 = torch.empty((ntot, 3), dtype=torch.float64, device=dev)
            c1.download_accel_device(a1.data_ptr())
            torch.cuda.synchronize()
            e = ((a0 - a1[mine]).norm(dim=1) / a1[mine].norm(dim=1).clamp(min=1e-300))
            verify = {"criterion": "Barnes-Hut pass of the bootstrap", "interactions_all_ranks": int(tot_bh), "interactions_one_rank": int(st1.interactions),
                      "equal": bool(int(tot_bh) == int(st1.interactions)), "rank0_rel_err_median": float(e.median()), "rank0_rel_err_max": float(e.max())}
            c1.close()
            del a1
            P.TreeUseBH = 0
            assert verify["equal"], verify
        dist.barrier()
    del base_pos, base_mass, a0
    pos_all = mass_all = None
    torch.cuda.empty_cache()

    def step():
        return D.step(pos, mass, oldacc)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    for _ in range(W):
        step()
    # one extra untimed step with CUDA-event marks between the phases: run sequentially on one stream so that the marks
    # mean something (the timed steps overlap the PM chain with the tree chain and carry no marks)
    D.profile = True
    D.stage_ms()
    step()
    phases = {k: round(v, 3) for k, v in D.stage_ms().items()}
    let = D.let_info
    D.profile = False
    step()
    sampler = ClockSampler(local_rank)
    sampler.start()
    B.loc.reset_timings(); B.tree.reset_timings(); B.pmc.reset_timings()
    D.comm.bytes_sent = 0
    if D.comm_pm is not None:
        D.comm_pm.bytes_sent = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    walk_ms, inter, entries, pseudo = [], [], [], 0
    for _ in range(args.steps):
        _, _, _, st = step()
        walk_ms.append(st.kernel_ms); inter.append(st.interactions); entries.append(st.list_entries); pseudo += st.pseudo_opened
    e1.record()
    barrier()
    dt = allmax(e0.elapsed_time(e1) * 1e-3)
    launches = B.loc.timings().kernel_launches + B.tree.timings().kernel_launches + B.pmc.timings().kernel_launches
    clocks = sampler.stop()
    sent = (D.comm.bytes_sent + (D.comm_pm.bytes_sent if D.comm_pm is not None else 0)) / args.steps + let.bytes_sent
    ms_per_step = 1e3 * dt / args.steps
    total_inter = allsum(float(np.mean(inter)))
    wms_mine = float(np.mean(walk_ms))
    wms_max, wms_min = allmax(wms_mine), allmin(wms_mine)
    achieved_tf = float(np.mean(inter)) * FLOP_PER_INTERACTION / (wms_mine * 1e-3) / 1e12     # rank 0's kernel against its own SMs
    assert allsum(float(pseudo)) == 0, "a pseudo node was opened without an import below it"
    mem = torch.cuda.mem_get_info(dev)
    mem_used_max = allmax(float(mem[1] - mem[0]))

    # end to end: host buffers in, host buffers out, copies inside the timed region
    de, e2e_steps = float("nan"), 1
    if not args.no_e2e:
        hpos, hmass, hold = pos.cpu().pin_memory(), mass.cpu().pin_memory(), oldacc.cpu().pin_memory()
        hacc = torch.empty((n, 3), dtype=torch.float64).pin_memory(); hpot = torch.empty(n, dtype=torch.float64).pin_memory()
        hpm = torch.empty((n, 4), dtype=torch.float64).pin_memory()

        def e2e_step():
            dp, dm, do = hpos.to(dev, non_blocking=True), hmass.to(dev, non_blocking=True), hold.to(dev, non_blocking=True)
            a, p, o, st = D.step(dp, dm, do, pm_host_out=hpm)
            hacc.copy_(a, non_blocking=True); hpot.copy_(p, non_blocking=True)
            torch.cuda.synchronize()

        e2e_step()
        barrier()
        t0 = time.perf_counter()
        e2e_steps = max(2, min(args.steps, 5))
        for _ in range(e2e_steps):
            e2e_step()
        barrier()
        de = allmax(time.perf_counter() - t0)
    ntot_all = int(allsum(float(n)))
    n_max, n_min = int(allmax(float(n))), int(allmin(float(n)))
    imp_p_max, imp_n_max = int(allmax(float(let.imported_particles))), int(allmax(float(let.imported_nodes)))

    workload_name = ("clustered-%d^3-box%gkpc-nmesh%d (%s)" % (ngrid, box, nmesh, ics_tag)) if clustered else "zeldovich-%d^3-nmesh%d-z9" % (ngrid, nmesh)
    line = {"metric": "short-range gravity particles/s", "value": ntot_all / (dt / args.steps), "unit": "particles/s", "n_gpus": world,
            "steps": args.steps, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 pair math, f64 accumulation/tests", "data": "synthetic",
            "config": {"workload": "%s domain-decomposed over %d GPUs (%.3f x 256^3 particles per GPU)" % (workload_name, world, ntot_all / world / 256. ** 3),
                       "particles": ntot_all, "particles_per_gpu_min_max": [n_min, n_max], "Nmesh": nmesh, "ErrTolForceAcc": 0.002, "TreeRcut": 6.0, "Asmth": 1.5,
                       "group_width": args.group_width,
                       "parallelism": "top tree over PH key ranges (%d top nodes, %d top leaves, %s per rank), own sub-trees + imported boundary sub-trees (raw NCCL in the library), x-slab PM on a second stream + communicator"
                                      % (len(dom.topnodes), dom.ntopleaves, "/".join(str(int(t["EndLeaf"] - t["StartLeaf"])) for t in dom.tasks)),
                       "balance": balance,
                       "l2": "inputs_exceed_l2 (particle arrays %.0f MB per GPU)" % (n * 36 / 1e6)},
            "interactions_per_s": total_inter / (wms_max * 1e-3), "interactions_per_particle": total_inter / ntot_all,
            "walk_particles_per_s": ntot_all / (wms_max * 1e-3),
            "stage_ms": {"walk_ms_max_over_ranks": wms_max, "walk_ms_min_over_ranks": wms_min, "walk_max_over_min": wms_max / wms_min, "walk_ms_rank0": wms_mine,
                         "other_ms (local sort, tree build, exchanges, PM)": ms_per_step - wms_max,
                         "phases_rank0_ms (one untimed step run sequentially, CUDA events)": phases,
                         "tree_build_dist_rank0_ms": {"local_build": let.build_ms, "moments_exchange": let.moments_exchange_ms, "mark+pack": let.let_mark_ms,
                                                      "import_exchange": let.let_exchange_ms, "attach": let.let_attach_ms}},
            "memory": {"per_gpu": "O(N/P + boundary)", "local_particles_max": n_max, "imported_particles_max": imp_p_max, "imported_nodes_max": imp_n_max,
                       "import_fraction_max": imp_p_max / float(max(1, n_max)), "local_nodes_rank0": int(let.local_nodes), "device_bytes_in_use_max": int(mem_used_max)},
            "exchange_bytes_sent_per_step_rank0": int(sent), "tree_nodes": int(D.tree_info.numnodes),
            "parity_vs_one_rank": verify,
            "roofline": {"bound": "fp32", "kernel": "sq_walk_kernel", "achieved": achieved_tf, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": achieved_tf / fp32_peak if fp32_peak > 0 else None, "traffic": None, "flop_per_interaction": FLOP_PER_INTERACTION,
                         "peak_source": "FFMA microbenchmark in this run (sq_measure_fp32_peak), rank 0",
                         "lane_efficiency": float(np.mean(inter)) / float(max(1, np.mean(entries) * 32))},
            "e2e": None if args.no_e2e else {"value": ntot_all / (de / e2e_steps), "unit": "particles/s", "h2d_bytes_per_step": int(ntot_all * (24 + 4 + 8)),
                    "d2h_bytes_per_step": int(ntot_all * (24 + 8 + 32)), "ms_per_step": 1e3 * de / e2e_steps},
            "gpu_launches": int(launches), "clocks": clocks}
    if rank == 0:
        print(json.dumps(line))
    B.close()
    dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------- GPU side
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--ngrid", type=int, default=256)
    ap.add_argument("--nmesh", type=int, default=0)
    ap.add_argument("--cpu-ngrid", type=int, default=64)
    ap.add_argument("--group-width", type=int, default=32)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--ics", default="zeldovich", choices=["zeldovich", "clustered"], help="N > 1: clustered = BASELINE config 3 (512^3 in a 50 Mpc/h box at N = 8)")
    ap.add_argument("--clustered-frac", type=float, default=0.5, help="fraction of the particles in Hernquist halos (shenqi_b200.ics.clustered)")
    ap.add_argument("--box", type=float, default=0.0, help="box size in kpc/h (default: 1 Mpc/h mean spacing; clustered: 50 Mpc/h x Ngrid / 512)")
    ap.add_argument("--balance", default="count", choices=["count", "cost"], help="N > 1: cut the top leaves into segments of equal particle count or equal measured interaction cost")
    ap.add_argument("--domain-level", type=int, default=5, help="N > 1: histogram level of the domain's top tree (8^level cells)")
    ap.add_argument("--overdecomposition", type=int, default=4, help="DomainOverDecompositionFactor (gadget/params.cpp:84)")
    ap.add_argument("--tile", type=int, default=1, help="N > 1: build the box from tile^3 periodic copies of an (ngrid / tile)^3 realisation (1024^3 = --ngrid 1024 --tile 2)")
    ap.add_argument("--deadline", type=float, default=420.0, help="N > 1: seconds after which every rank is ended by SIGALRM whatever it is waiting for")
    ap.add_argument("--no-e2e", action="store_true", help="N > 1: skip the host-buffer leg (e.g. 1024^3: 13 GB of pinned memory per rank)")
    ap.add_argument("--no-verify", dest="verify", action="store_false", help="N > 1: skip the one-rank parity run on rank 0")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    nmesh = args.nmesh or 2 * args.ngrid

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        return run_dist(args, rank, local_rank, world)
    import torch
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    import shenqi_b200 as sq

    W = max(args.warmup, 3)
    pos, mass, par = workload(args.ngrid, nmesh)
    n = len(pos)
    ctx = sq.Context(local_rank)
    ctx.set_group_width(args.group_width)
    P = sq.Params.defaults(par["BoxSize"], nmesh, par["GravitySoftening"], G=par["G"], rho0=par["rho0"], TreeRcut=par["TreeRcut"], TreeUseBH=1)
    ctx.set_params(P)
    fp32_peak = ctx.measure_fp32_peak()
    hbm_peak, peak_src = read_peaks()

    # bootstrap |a_old|: Barnes-Hut pass + PM, exactly as the reference starts (gravshort-tree2.cpp:170-171)
    ctx.upload_particles(pos, mass)
    ctx.ph_sort(); ctx.tree_build()
    a0, _, _ = ctx.grav_short_tree()
    g0, _ = ctx.gravpm_force()
    P.TreeUseBH = 0
    ctx.set_params(P)
    ctx.set_old_accel(a0, g0)

    def device_step():
        ctx.ph_sort()
        ctx.tree_build()
        st = ctx.grav_short_tree_device()
        ctx.gravpm_force_device()
        return st

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(W):
        device_step()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ctx.reset_timings()
    barrier()
    t0 = time.perf_counter()
    stats, walk_ms, stage = [], [], []
    for _ in range(args.steps):
        st = device_step()
        stats.append(st); walk_ms.append(st.kernel_ms)
        tm = ctx.timings()
        stage.append([getattr(tm, f) for f, _ in tm._fields_[:10]])
    barrier()
    dt = time.perf_counter() - t0
    launches = ctx.timings().kernel_launches
    clocks = sampler.stop()
    if dist is not None:
        tt = torch.tensor([dt], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
    ms_per_step = 1e3 * dt / args.steps
    value = world * n / (dt / args.steps)
    interactions = int(np.mean([s.interactions for s in stats]))
    wms = float(np.mean(walk_ms))
    achieved_tf = interactions * FLOP_PER_INTERACTION / (wms * 1e-3) / 1e12
    stage = np.mean(np.array(stage), axis=0)
    stage_names = ["keys_ms", "sort_ms", "gather_ms", "tree_build_ms", "tree_moments_ms", "walk_ms", "pm_paint_ms", "pm_fft_ms", "pm_transfer_ms", "pm_readout_ms"]
    stage_ms = dict(zip(stage_names, [float(x) for x in stage]))

    # HBM-bound stages: algorithmic bytes / device time / measured copy bandwidth
    numnodes = ctx.tree_info.numnodes
    nm3 = float(nmesh) ** 3
    bytes_model = {"keys": 32.0 * n, "sort": 8 * 32.0 * n, "gather": 2 * 36.0 * n + 8.0 * n,
                   "tree_build": 8.0 * n + 28.0 * n + 80.0 * numnodes, "pm_paint": 28.0 * n + 16.0 * nm3,
                   "pm_readout": 24.0 * n + 32.0 * n + 4 * 8.0 * nm3,
                   "pm_transfer": 32.0 * (nm3 / 2),                      # rho_k -> phi_k in place: 16 B read + 16 B written per mode
                   "pm_fft_gradient": (16.0 + 16.0 + 32.0) * nm3}          # D2Z + Z2D (8 B in, 8 B out each, ideal) + 4-point gradient (8 B in, 24 B out)
    hbm = {}
    for k, ms_key in (("keys", "keys_ms"), ("sort", "sort_ms"), ("gather", "gather_ms"), ("tree_build", "tree_build_ms"),
                      ("pm_paint", "pm_paint_ms"), ("pm_readout", "pm_readout_ms"), ("pm_transfer", "pm_transfer_ms"), ("pm_fft_gradient", "pm_fft_ms")):
        if stage_ms[ms_key] > 0:
            gbs = bytes_model[k] / (stage_ms[ms_key] * 1e-3) / 1e9
            hbm[k] = {"achieved_gbs": gbs, "frac": gbs / hbm_peak, "ms": stage_ms[ms_key]}

    # end to end through the C ABI with host buffers (pinned)
    ppos, _k1 = pinned(pos); pmass, _k2 = pinned(mass)
    pa0, _k3 = pinned(a0); pg0, _k4 = pinned(g0)
    acc_out, _k5 = pinned(np.zeros((n, 3))); pot_out, _k6 = pinned(np.zeros(n))
    gpm_out, _k7 = pinned(np.zeros((n, 3))); gpot_out, _k8 = pinned(np.zeros(n))

    def e2e_step():
        # one call of the public API with host buffers: sq_gravity_step (uploads, sort, tree, PM, walk, downloads)
        ctx.gravity_step(ppos, pmass, pa0, pg0, accel=acc_out, potential=pot_out, gravpm=gpm_out, pm_potential=gpot_out)

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(2, min(args.steps, 5))
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    de = time.perf_counter() - t0
    if dist is not None:
        tt = torch.tensor([de], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        de = float(tt.item())
    e2e_value = world * n / (de / e2e_steps)
    h2d = n * (24 + 4 + 48)                        # pos, mass, FullTreeGravAccel + GravPM (the |a_old| inputs)
    d2h = n * (24 + 8 + 24 + 8)

    line = {"metric": "short-range gravity particles/s", "value": value, "unit": "particles/s", "n_gpus": world, "steps": args.steps,
            "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 pair math, f64 accumulation/tests", "data": "synthetic",
            "config": bench_config(args.ngrid, nmesh, n, args.group_width),
            "interactions_per_s": world * interactions / (wms * 1e-3), "interactions_per_particle": interactions / n,
            "walk_particles_per_s": world * n / (wms * 1e-3),
            "stage_ms": stage_ms, "tree_nodes": int(numnodes),
            "roofline": {"bound": "fp32", "kernel": "sq_walk_kernel", "achieved": achieved_tf, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": achieved_tf / fp32_peak if fp32_peak > 0 else None, "traffic": WALK_DRAM_TRAFFIC_256 if args.ngrid == 256 else None, "traffic_unit": "bytes per launch (ncu dram read + write)",
                         "flop_per_interaction": FLOP_PER_INTERACTION, "peak_source": "FFMA microbenchmark in this run (sq_measure_fp32_peak)",
                         "lane_efficiency": interactions / float(max(1, np.mean([s.list_entries for s in stats]) * 32))},
            "roofline_hbm": {"peak_gbs": hbm_peak, "peak_source": peak_src, "stages": hbm},
            "e2e": {"value": e2e_value, "unit": "particles/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * de / e2e_steps},
            "gpu_launches": int(launches), "clocks": clocks}

    if rank == 0 and not args.no_cpu:
        # CPU baseline on a bounded sample of the same workload, and the SAME sample through the GPU path: the two sets of
        # accelerations are diffed (median / p99.9 of |a - a_ref| / |a_ref|, interaction counts) and reported in the line
        cng = args.cpu_ngrid
        cp, cm, cpar = workload(cng, cng * nmesh // args.ngrid)
        Pc = sq.Params.defaults(cpar["BoxSize"], cpar["Nmesh"], cpar["GravitySoftening"], G=cpar["G"], rho0=cpar["rho0"], TreeRcut=cpar["TreeRcut"], TreeUseBH=1)
        ctx.set_params(Pc)
        ctx.upload_particles(cp, cm); ctx.ph_sort(); ctx.tree_build()
        b0, _, _ = ctx.grav_short_tree()
        h0, _ = ctx.gravpm_force()
        Pc.TreeUseBH = 0
        ctx.set_params(Pc)
        ctx.set_old_accel(b0, h0)
        b1, _, stc = ctx.grav_short_tree()
        os.environ["OMP_NUM_THREADS"] = str(cpu_threads())
        res = cpu_reference_step(cp, cm, cpar, old=(b0, h0))
        err = np.linalg.norm(b1 - res["accel"], axis=1) / np.maximum(np.linalg.norm(res["accel"], axis=1), 1e-300)
        line["cpu_baseline"] = {"value": len(cp) / res["step_s"], "unit": "particles/s", "cores": res["cores"], "kind": res["kind"],
                                "sample": "periodic %d^3 / Nmesh %d sub-volume of the workload (same spacing/softening/Rcut/ErrTol): sort+tree+walk by the %s; PM restatement timed apart" % (cng, cpar["Nmesh"], "reference's own OpenMP code (oracle/_ref)" if res["kind"] == "reference" else "C port (oracle/)"),
                                "interactions_per_s": res["interactions"] / res["walk_s"], "seconds": {k: res[k] for k in ("sort_s", "tree_s", "walk_s", "pm_s", "pm_ref_s")},
                                "parity_on_sample": {"gpu_interactions": int(stc.interactions), "cpu_interactions": int(res["interactions"]),
                                                     "interactions_equal": bool(int(stc.interactions) == int(res["interactions"])),
                                                     "accel_rel_err_median": float(np.median(err)), "accel_rel_err_p999": float(np.percentile(err, 99.9))}}
        if res.get("gravpm_ref") is not None:
            # the GPU PM against gravpm_force of the reference itself (its own gravpm.cpp + petapm.cpp; stand-in FFT)
            gr = res["gravpm_ref"]
            line["cpu_baseline"]["parity_on_sample"]["pm_accel_max_abs_err_over_max"] = float(np.abs(h0 - gr).max() / np.abs(gr).max())
    if rank == 0:
        print(json.dumps(line))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
