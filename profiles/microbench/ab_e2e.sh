# A/B of libstitch builds on the e2e leg (host buffers): usage  bash profiles/microbench/ab_e2e.sh lib1.so lib2.so ...
for rep in 1 2; do for lib in "$@"; do
  STITCH_LIB=$lib python bench.py --steps 100 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('$lib', round(d['ms_per_step']*1e3,2), 'e2e ms', round(d['e2e']['ms_per_step'],3), 'single', round(d['e2e']['single_job_latency_ms'],3))"
done; done
