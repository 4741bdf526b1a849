# schedule sweep of the upscale units on BASELINE configs[0]: body fraction x tail unit height
for f in ${FRACS:-0.8 0.9 0.95 1.0}; do for t in ${TAILS:-8 16}; do
  r=$(STITCH_UP_BODY_FRAC=$f STITCH_UP_TAIL_H=$t python bench.py --workload cfg1 --lanes 1 --steps 200 --warmup 20 --no-cpu-baseline --e2e-steps 3 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(round(d['ms_per_step']*1e3,2))")
  echo "frac=$f tail=$t us=$r"
done; done
