# guided-schedule knobs of the streaming regions (launch.cuh) for the current kernel, cfg2 + cfg2_yuv
run() { python bench.py --workload $1 --no-cpu-baseline --e2e-steps 3 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('%-10s %-40s %7.2f us frac %.3f' % ('$1', '$2', d['ms_per_step']*1e3, d['roofline']['frac']))"; }
for w in cfg2 cfg2_yuv; do
  run $w "default"
  for bf in 0.6 0.85 0.95; do STITCH_BODY_FRAC=$bf run $w "BODY_FRAC=$bf"; done
  for tr in 16 64 96; do STITCH_TAIL_ROWS=$tr run $w "TAIL_ROWS=$tr"; done
  STITCH_BODY_FRAC=0.85 STITCH_TAIL_ROWS=64 run $w "BODY_FRAC=0.85 TAIL_ROWS=64"
  STITCH_LIB=$PWD/variants/libstitch_evict.so run $w "evict_first"
  python bench.py --workload $w --lanes 1 --no-cpu-baseline --e2e-steps 3 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('%-10s %-40s %7.2f us frac %.3f' % ('$w', 'lanes=1', d['ms_per_step']*1e3, d['roofline']['frac']))"
  python bench.py --workload $w --lanes 3 --no-cpu-baseline --e2e-steps 3 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('%-10s %-40s %7.2f us frac %.3f' % ('$w', 'lanes=3', d['ms_per_step']*1e3, d['roofline']['frac']))"
done
