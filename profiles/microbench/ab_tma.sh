# A/B of the streaming unit's staging on one box: TMA ring (default build) vs the variants under variants/
# (make -C imagestitching_b200/csrc OUT=... BUILD=... EXTRA_NVFLAGS=-DSTITCH_STREAM_TMA=0 etc.)
for lib in default $(ls variants/*.so 2>/dev/null); do
  for w in cfg2 cfg2_yuv cfg3 cfg3_min cfg2_android; do
    if [ $lib = default ]; then unset STITCH_LIB; else export STITCH_LIB=$PWD/$lib; fi
    timeout 300 python bench.py --workload $w --no-cpu-baseline --e2e-steps 3 2>/dev/null | python -c "
import sys,json
ls=[l for l in sys.stdin if l.startswith('{')]
if not ls: print('$lib $w FAILED'); sys.exit(0)
d=json.loads(ls[-1]); r=d['roofline']; c=d['clocks']
print('%-34s %-12s %7.2f us  %6.0f GB/s  frac %.3f  SM %s MHz %s' % ('$lib', '$w', d['ms_per_step']*1e3, r['achieved'], r['frac'], c['sm_mhz'], c['reasons']))"
  done
done
