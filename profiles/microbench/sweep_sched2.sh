run() { python bench.py --workload $1 --no-cpu-baseline --e2e-steps 3 $3 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('%-12s %-44s %7.2f us frac %.3f %s' % ('$1', '$2 $3', d['ms_per_step']*1e3, d['roofline']['frac'], d['clocks']['sm_mhz']))"; }
for cfg in "0.75 96" "0.75 128" "0.75 160" "0.75 216" "0.9 96" "0.9 128" "0.9 216" "1.0 32" "0.5 128" "0.6 216"; do
  set -- $cfg
  export STITCH_BODY_FRAC=$1 STITCH_TAIL_ROWS=$2
  run cfg2 "bf=$1 tr=$2"
  run cfg2 "bf=$1 tr=$2" "--lanes 1"
  run cfg2_yuv "bf=$1 tr=$2"
done
for cfg in "0.75 32" "0.75 96" "0.75 128" "0.9 128"; do
  set -- $cfg
  export STITCH_BODY_FRAC=$1 STITCH_TAIL_ROWS=$2
  for w in cfg3 cfg3_min cfg2_android cfg5; do run $w "bf=$1 tr=$2"; done
done
