# A/B of libstitch builds on the JPEG export kernels: usage  bash profiles/microbench/ab_jpeg.sh lib1.so lib2.so ...
for lib in "$@"; do
  STITCH_LIB=$lib ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file /tmp/ab_jpeg.csv python bench.py --workload cfg2_jpeg --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 3 > /dev/null 2>&1
  echo "== $lib"; python profiles/launch_summary.py /tmp/ab_jpeg.csv | grep jpeg
  r=$(STITCH_LIB=$lib python bench.py --workload cfg2_jpeg --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 3 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(round(d['ms_per_step']*1e3,2))")
  echo "job us=$r"
done
