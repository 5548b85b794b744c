#!/bin/bash
# BASELINE config 4: 1024^3 particles (2^3 tiles of a 512^3 Zel'dovich box), Nmesh 2048, 8 GPUs
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
free -g | head -2
timeout 900 $TR 29524 bench.py --gpus 8 --steps 3 --warmup 3 --ngrid 1024 --tile 2 --no-e2e --no-verify > gpurun_out/bench_r2_n8_c4.json 2> gpurun_out/bench_r2_n8_c4.err; grep -v "^\[W\|^W1" gpurun_out/bench_r2_n8_c4.err | tail -c 1500
python - <<'PY'
import json
for f in ("n8_c4",):
    try:
        d=json.loads(open("gpurun_out/bench_r2_%s.json"%f).read().strip().splitlines()[-1])
        s=d["stage_ms"]
        print(f, "%.1fM p/s"%(d["value"]/1e6), "%.1f ms"%d["ms_per_step"], "walk max/min %.1f/%.1f"%(s["walk_ms_max_over_ranks"], s["walk_ms_min_over_ranks"]), s["tree_build_dist_rank0_ms"], d["memory"], d["config"]["particles_per_gpu_min_max"], d["interactions_per_particle"])
        print("   ", s["phases_rank0_ms (one untimed step run sequentially, CUDA events)"])
    except Exception as e: print(f, "ERR", e)
PY
nvidia-smi --query-gpu=memory.used --format=csv | head -3
