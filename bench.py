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


def cpu_reference_step(pos, mass, par, old=None, want="auto", with_pm=True):
    """One step of the reference's own CPU code (oracle/_ref: slots_gc_sorted, force_tree_full, grav_short_tree, OpenMP, all
    host cores) or, if that was not built, of the C port (oracle/).  old = (FullTreeGravAccel, GravPM) inputs of the
    relative criterion in input order (None: bootstrapped by an untimed Barnes-Hut pass).  petapm.cpp needs FFTW/heFFTe
    and cannot be built here: the PM part is the numpy restatement, timed separately (pm_s) and NOT part of step_s.
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
    out["pm_s"] = None
    if with_pm:
        t0 = time.perf_counter(); H.o_pm_force(pos, mass, box, par["Nmesh"], 1.5, par["G"]); out["pm_s"] = time.perf_counter() - t0
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
    pm = cpu_reference_step(pos, mass, par, with_pm=True)["pm_s"] if ngrid <= 128 else None
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
                             "seconds": {k: res[k] for k in ("sort_s", "tree_s", "walk_s")}, "pm_restatement_s (numpy, not in value)": pm},
            "e2e": {"value": v, "unit": "particles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------- multi-GPU
def run_dist(args, rank, local_rank, world):
    """N > 1: ONE domain-decomposed problem (not replicas): ~256^3 particles per GPU of a global Zel'dovich box, every
    rank owning a consecutive range of the global Peano-Hilbert order (what the reference's domain decomposition hands
    the path).  Step = shenqi_b200.dist.DistGravity: local sort, all-gather of the sorted ranges, global tree, walk of
    the rank's own targets; PM on x-slabs with all-to-all routing/transposes.  NCCL over NVLink."""
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":              # NCCL prints its banner on stdout: keep it to the one JSON line
        del os.environ["NCCL_DEBUG"]
    dist.init_process_group("nccl", device_id=dev)
    import shenqi_b200 as sq
    from shenqi_b200.ics import zeldovich_torch, rho0, G_INTERNAL
    from shenqi_b200.dist import DistGravity, CudaBackend, Comm

    ngrid = args.ngrid if args.ngrid != 256 else DIST_NGRID.get(world, int(round(256 * world ** (1 / 3.) / 8)) * 8)
    nmesh = args.nmesh or 2 * ngrid
    W = max(args.warmup, 3)
    box = 1000.0 * ngrid
    P = sq.Params.defaults(box, nmesh, (box / ngrid) / 30., G=G_INTERNAL, rho0=rho0(), TreeRcut=6.0, TreeUseBH=1)
    B = CudaBackend(P, local_rank)
    B.glob.set_group_width(args.group_width)
    # the PM exchanges get their own communicator so that DistGravity.step() can run them underneath the tree's
    D = DistGravity(B, Comm(), profile=False, comm_pm=Comm(dist.new_group()))
    fp32_peak = B.glob.measure_fp32_peak()
    hbm_peak, peak_src = read_peaks()

    # ---- set-up (untimed): global ICs on every GPU, domain = equal-count range of the global PH order, shuffled
    pos_all, mass_all = zeldovich_torch(ngrid, box, dev)
    ntot = pos_all.shape[0]
    srt = B.sort_local(pos_all, mass_all)
    lo, hi = rank * ntot // world, (rank + 1) * ntot // world
    g = torch.Generator(device=dev); g.manual_seed(1234 + rank)
    mine = srt["perm"][lo:hi][torch.randperm(hi - lo, device=dev, generator=g)]
    pos = pos_all[mine].contiguous(); mass = mass_all[mine].contiguous()
    del pos_all, mass_all, srt, mine
    n = pos.shape[0]
    torch.cuda.empty_cache()

    # bootstrap |a_old| with a Barnes-Hut pass + PM, as the reference starts (gravshort-tree2.cpp:170-171)
    a0, _, _ = D.tree_force(pos, mass)
    o4 = D.pm_force(pos, mass)
    oldacc = ((a0 + o4[:, 1:4]).norm(dim=1) / G_INTERNAL).contiguous()
    del a0, o4
    P.TreeUseBH = 0
    B.set_params(P)

    def step():
        return D.step(pos, mass, oldacc)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    for _ in range(W):
        step()
    # one extra untimed step with CUDA-event marks between the phases: run sequentially on one stream so that the marks
    # mean something (the timed steps overlap the PM chain with the tree chain and carry no marks)
    D.profile = True
    D.stage_ms()
    step()
    phases = {k: round(v, 3) for k, v in D.stage_ms().items()}
    D.profile = False
    step()
    sampler = ClockSampler(local_rank)
    sampler.start()
    B.loc.reset_timings(); B.glob.reset_timings()
    D.comm.bytes_sent = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    walk_ms, inter, entries, tb_ms = [], [], [], []
    for _ in range(args.steps):
        _, _, _, st = step()
        walk_ms.append(st.kernel_ms); inter.append(st.interactions); entries.append(st.list_entries)
        tg = B.glob.timings()
        tb_ms.append(tg.tree_build_ms + tg.tree_moments_ms)
    e1.record()
    barrier()
    dt = allmax(e0.elapsed_time(e1) * 1e-3)
    launches = B.loc.timings().kernel_launches + B.glob.timings().kernel_launches
    clocks = sampler.stop()
    sent = D.comm.bytes_sent / args.steps
    ms_per_step = 1e3 * dt / args.steps
    total_inter = allsum(float(np.mean(inter)))
    wms_max = allmax(float(np.mean(walk_ms)))
    wms_mine = float(np.mean(walk_ms))
    achieved_tf = float(np.mean(inter)) * FLOP_PER_INTERACTION / (wms_mine * 1e-3) / 1e12     # rank 0's kernel against its own SMs
    numnodes = int(D.tree_info.numnodes)

    # end to end: host buffers in, host buffers out, copies inside the timed region
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

    line = {"metric": "short-range gravity particles/s", "value": ntot_all / (dt / args.steps), "unit": "particles/s", "n_gpus": world,
            "steps": args.steps, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 pair math, f64 accumulation/tests", "data": "synthetic",
            "config": {"workload": "zeldovich-%d^3-nmesh%d-z9 domain-decomposed over %d GPUs (%.3f x 256^3 particles per GPU)" % (ngrid, nmesh, world, n / 256. ** 3),
                       "particles": ntot_all, "particles_per_gpu": n, "Nmesh": nmesh, "ErrTolForceAcc": 0.002, "TreeRcut": 6.0, "Asmth": 1.5,
                       "group_width": args.group_width, "parallelism": "PH-range domains, replicated global tree, x-slab PM (PM chain on a second stream + communicator underneath the tree's all-gather and build)",
                       "l2": "inputs_exceed_l2 (particle arrays %.0f MB per GPU)" % (ntot_all * 36 / 1e6)},
            "interactions_per_s": total_inter / (wms_max * 1e-3), "interactions_per_particle": total_inter / ntot_all,
            "walk_particles_per_s": ntot_all / (wms_max * 1e-3),
            "stage_ms": {"walk_ms_max_over_ranks": wms_max, "walk_ms_rank0": wms_mine, "global_tree_build_ms_rank0": float(np.mean(tb_ms)),
                         "other_ms (local sort, exchanges, PM)": ms_per_step - wms_max - float(np.mean(tb_ms)),
                         "phases_rank0_ms (one untimed step run sequentially, CUDA events)": phases},
            "exchange_bytes_sent_per_step_rank0": int(sent), "tree_nodes": numnodes,
            "roofline": {"bound": "fp32", "kernel": "sq_walk_kernel", "achieved": achieved_tf, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": achieved_tf / fp32_peak if fp32_peak > 0 else None, "traffic": None, "flop_per_interaction": FLOP_PER_INTERACTION,
                         "peak_source": "FFMA microbenchmark in this run (sq_measure_fp32_peak), rank 0",
                         "lane_efficiency": float(np.mean(inter)) / float(max(1, np.mean(entries) * 32))},
            "e2e": {"value": ntot_all / (de / e2e_steps), "unit": "particles/s", "h2d_bytes_per_step": int(n * (24 + 4 + 8)) * world,
                    "d2h_bytes_per_step": int(n * (24 + 8 + 32)) * world, "ms_per_step": 1e3 * de / e2e_steps},
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
                                "interactions_per_s": res["interactions"] / res["walk_s"], "seconds": {k: res[k] for k in ("sort_s", "tree_s", "walk_s", "pm_s")},
                                "parity_on_sample": {"gpu_interactions": int(stc.interactions), "cpu_interactions": int(res["interactions"]),
                                                     "interactions_equal": bool(int(stc.interactions) == int(res["interactions"])),
                                                     "accel_rel_err_median": float(np.median(err)), "accel_rel_err_p999": float(np.percentile(err, 99.9))}}
    if rank == 0:
        print(json.dumps(line))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
