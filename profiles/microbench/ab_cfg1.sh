# A/B of libstitch builds on the upscale workloads: usage  bash profiles/microbench/ab_cfg1.sh lib1.so lib2.so ...
for lib in "$@"; do for wl in cfg1 cfg3; do
  r=$(STITCH_LIB=$lib python bench.py --workload $wl --lanes $([ $wl = cfg1 ] && echo 1 || echo 2) --steps 200 --warmup 20 --no-cpu-baseline --e2e-steps 3 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(round(d['ms_per_step']*1e3,2))")
  echo "$lib $wl us=$r"
done; done
