#!/usr/bin/env python
"""Times the SURVEY 8(f2) kernels (drift, kick, active sub-list) at the 256^3 load of BASELINE config 2 with CUDA events
on the launching stream and prints one JSON line: ms per call and achieved GB/s of the ALGORITHMIC bytes
(drift 72 B/particle, kick 72 B/active particle (+4 B index with a list), sub-list 1 B timebin x 2 passes + 4 B per
selected entry (+ 4 B x 2 per candidate with a list)) against MEASURED_PEAKS.json:hbm_gbs.  Inputs (1.2 GB) exceed L2."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import shenqi_b200 as sq


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    n, box = 256 ** 3, 100000.0
    peak = 6459.6
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    ctx = sq.Context(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cuda").manual_seed(1)
    pos = torch.rand((n, 3), dtype=torch.float64, device="cuda", generator=g) * box + 1e-9
    vel = torch.randn((n, 3), dtype=torch.float64, device="cuda", generator=g) * 100.0
    acc = torch.randn((n, 3), dtype=torch.float64, device="cuda", generator=g)
    timebin = torch.randint(1, 9, (n,), dtype=torch.uint8, device="cuda", generator=g)
    out = torch.empty(n, dtype=torch.int32, device="cuda")
    res = {"particles": n, "hbm_peak_gbs": peak, "stages": {}}

    def put(name, ms, nbytes, **kw):
        res["stages"][name] = dict(ms=ms, algorithmic_bytes=nbytes, achieved_gbs=nbytes / ms / 1e6, frac=nbytes / ms / 1e6 / peak, **kw)

    ms = timed(lambda: ctx.drift_device(pos.data_ptr(), vel.data_ptr(), n, 0.01, [1.0, 2.0, 3.0], box))
    put("drift", ms, 72 * n, note="includes the 8-byte error-counter read-back and stream sync of each call")
    ms = timed(lambda: ctx.kick_device(vel.data_ptr(), acc.data_ptr(), n, 1e-3))
    put("kick_all", ms, 72 * n)
    for ti, label in ((1 << 8, "all_bins_active"), (1 << 3, "bins<=3"), (1 << 1, "bin1")):
        m = ctx.build_active_sublist_device(n, timebin.data_ptr(), 46, ti, out.data_ptr())
        ms = timed(lambda: ctx.build_active_sublist_device(n, timebin.data_ptr(), 46, ti, out.data_ptr()))
        put("sublist_" + label, ms, 2 * n + 4 * m, selected=m)
        sub = out[:m].clone()
        ms = timed(lambda: ctx.kick_device(vel.data_ptr(), acc.data_ptr(), m, 1e-3, active_ptr=sub.data_ptr()))
        put("kick_" + label, ms, 76 * m, selected=m, note="gathers 24-B rows through the list: 2 sectors per row when sparse")
        out2 = torch.empty(max(m, 1), dtype=torch.int32, device="cuda")
        ms = timed(lambda: ctx.build_active_sublist_device(m, timebin.data_ptr(), 2, ti, out2.data_ptr(), active_ptr=sub.data_ptr()))
        m2 = ctx.build_active_sublist_device(m, timebin.data_ptr(), 2, ti, out2.data_ptr(), active_ptr=sub.data_ptr())
        put("sublist_of_" + label, ms, 10 * m + 4 * m2, candidates=m, selected=m2)
    import math
    pm = torch.randn((n, 3), dtype=torch.float64, device="cuda", generator=g)
    P = sq.TimebinParams(atime=0.1, hubble=3.2, ErrTolIntAccuracy=0.02, MinSizeTimestep=0.0, softening=1.0 / 30, Ti_Current=0, dti_max=1 << 40,
                         loga=math.log(0.1), lastsnap=0, can_advance=1, largest_active=40)
    P.sync_loga[0], P.sync_loga[1], P.sync_loga[2] = math.log(0.1), math.log(0.111), math.log(0.125)
    tb = torch.empty(n, dtype=torch.uint8, device="cuda")
    ms = timed(lambda: ctx.gravity_timebins_device(P, acc.data_ptr(), pm.data_ptr(), n, tb.data_ptr()))
    put("gravity_timebins", ms, 49 * n, note="48 B accelerations read + 1 B bin written per particle; includes the read-back of the 47 counts and the stream sync")
    ms = timed(lambda: ctx.timebin_pushdown_device(n, tb.data_ptr(), 20))
    put("timebin_pushdown", ms, 1 * n, note="1 B read per particle, writes only where the bin changes")
    print(json.dumps(res))
    ctx.close()


if __name__ == "__main__":
    main()
