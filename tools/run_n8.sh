#!/bin/bash
# one 8-GPU session: weak-scaling point, BASELINE config 3 (clustered 512^3, 50 Mpc/h, Nmesh 1024, cost-balanced), config 4 (1024^3, Nmesh 2048)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
nvidia-smi -L | wc -l; free -g | head -2
timeout 400 $TR 29521 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_r2_n8_zel.json 2> gpurun_out/bench_r2_n8_zel.err; tail -c 600 gpurun_out/bench_r2_n8_zel.err
timeout 400 $TR 29522 bench.py --gpus 8 --steps 5 --warmup 3 --ics clustered --balance cost > gpurun_out/bench_r2_n8_c3.json 2> gpurun_out/bench_r2_n8_c3.err; tail -c 600 gpurun_out/bench_r2_n8_c3.err
timeout 400 $TR 29523 bench.py --gpus 8 --steps 5 --warmup 3 --ics clustered --balance count --no-e2e > gpurun_out/bench_r2_n8_c3_count.json 2> gpurun_out/bench_r2_n8_c3_count.err; tail -c 600 gpurun_out/bench_r2_n8_c3_count.err
timeout 600 $TR 29524 bench.py --gpus 8 --steps 3 --warmup 3 --ngrid 1024 --tile 2 --no-e2e --no-verify > gpurun_out/bench_r2_n8_c4.json 2> gpurun_out/bench_r2_n8_c4.err; tail -c 1200 gpurun_out/bench_r2_n8_c4.err
python - <<'PY'
import json
for f in ("n8_zel","n8_c3","n8_c3_count","n8_c4"):
    try:
        d=json.loads(open("gpurun_out/bench_r2_%s.json"%f).read().strip().splitlines()[-1])
        s=d["stage_ms"]
        print(f, "%.1fM p/s"%(d["value"]/1e6), "%.1f ms"%d["ms_per_step"], "e2e", d["e2e"] and "%.1f ms"%d["e2e"]["ms_per_step"], "walk max/min %.1f/%.1f"%(s["walk_ms_max_over_ranks"], s["walk_ms_min_over_ranks"]), s["tree_build_dist_rank0_ms"], d["parity_vs_one_rank"], d["memory"], d["config"]["particles_per_gpu_min_max"], d["interactions_per_particle"])
        print("   ", s["phases_rank0_ms (one untimed step run sequentially, CUDA events)"])
    except Exception as e: print(f, "ERR", e)
PY
