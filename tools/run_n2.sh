TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
$TR 29511 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/bench_r2_n2_b.json 2> gpurun_out/bench_r2_n2_b.err; tail -c 800 gpurun_out/bench_r2_n2_b.err
$TR 29512 bench.py --gpus 2 --steps 3 --warmup 3 --ngrid 288 --ics clustered --balance cost > gpurun_out/bench_r2_n2_clu.json 2> gpurun_out/bench_r2_n2_clu.err; tail -c 800 gpurun_out/bench_r2_n2_clu.err
$TR 29513 bench.py --gpus 2 --steps 3 --warmup 3 --ngrid 288 --ics clustered --balance count > gpurun_out/bench_r2_n2_clu_count.json 2> gpurun_out/bench_r2_n2_clu_count.err; tail -c 800 gpurun_out/bench_r2_n2_clu_count.err
$TR 29514 bench.py --gpus 2 --steps 3 --warmup 3 --ngrid 288 --tile 2 > gpurun_out/bench_r2_n2_tile.json 2> gpurun_out/bench_r2_n2_tile.err; tail -c 800 gpurun_out/bench_r2_n2_tile.err
python - <<'PY'
import json
for f in ("n2_b","n2_clu","n2_clu_count","n2_tile"):
    try:
        d=json.loads(open("gpurun_out/bench_r2_%s.json"%f).read().strip().splitlines()[-1])
        s=d["stage_ms"]
        print(f, "%.1fM p/s"%(d["value"]/1e6), "%.1f ms"%d["ms_per_step"], "e2e %.1f ms"%d["e2e"]["ms_per_step"], "walk max/min %.1f/%.1f"%(s["walk_ms_max_over_ranks"], s["walk_ms_min_over_ranks"]), s["tree_build_dist_rank0_ms"], d["parity_vs_one_rank"], d["memory"]["import_fraction_max"], d["config"]["particles_per_gpu_min_max"], d["interactions_per_particle"], d["config"]["parallelism"][:80])
    except Exception as e: print(f, "ERR", e)
PY
