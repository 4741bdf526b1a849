# N-GPU record of one box: bash profiles/multi_gpu.sh N   (under gpurun --gpus N)
# (a) bench.py default (cfg2: device-resident, e2e with RGBA host buffers, e2e_files) as N torchrun ranks
# (b) BASELINE configs[4] (10 000 jobs sharded over the ranks, digests compared across ranks)
# (c) the same sweep in ONE process through stitch_pool_* (worker thread + ctx per GPU)
N=${1:-1}
run() { if [ $N -gt 1 ]; then python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; else python bench.py --gpus 1 "$@"; fi; }
run --no-cpu-baseline 2>/dev/null | grep '^{' | tail -1 > gpurun_out/r2_scale_cfg2_N$N.json
run --workload cfg5 --no-cpu-baseline --files-steps 0 2>/dev/null | grep '^{' | tail -1 > gpurun_out/r2_scale_cfg5_N$N.json
python bench.py --pool $N 2>/dev/null | grep '^{' | tail -1 > gpurun_out/r2_scale_pool_N$N.json
python - <<PY
import json
for k in ("cfg2", "cfg5", "pool"):
    try:
        d = json.load(open("gpurun_out/r2_scale_%s_N$N.json" % k))
    except Exception as e:
        print(k, "FAILED", e); continue
    e = d.get("e2e", {}); f = d.get("e2e_files", {})
    print("N=$N %-5s value %9.0f MP/s  %.1f us/job/gpu  frac %.3f  SM %s %s | e2e %s MP/s (%s ms) | files png %s MP/s (%s ms) jpeg %s MP/s (%s ms) | digests %s" % (
        k, d["value"], d["ms_per_step"] * 1e3 * (d["n_gpus"] if k == "pool" else 1), d["roofline"]["frac"], d["clocks"]["sm_mhz"], d["clocks"]["reasons"],
        round(e.get("value", 0)), round(e.get("ms_per_step", 0), 2), round(f.get("png", {}).get("value", 0)), round(f.get("png", {}).get("ms_per_step", 0), 2),
        round(f.get("jpeg", {}).get("value", 0)), round(f.get("jpeg", {}).get("ms_per_step", 0), 2), d.get("digests")))
PY
