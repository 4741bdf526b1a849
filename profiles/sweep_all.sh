# every bench workload on one box, one markdown row each:  bash profiles/sweep_all.sh > gpurun_out/sweep.md
echo "| workload | canvas | us/job | output MP/s | GB/s (algorithmic) | of HBM copy peak | e2e ms/job | e2e MP/s | SM MHz | host us/submit |"
echo "|---|---|---|---|---|---|---|---|---|---|"
for w in cfg2 cfg2_yuv cfg2_android cfg3 cfg3_min cfg1 cfg2_jpeg cfg2_png cfg5; do
  extra=""; [ $w = cfg2_png ] && extra="--steps 100"; [ $w = cfg2_jpeg ] && extra="--steps 200"
  python bench.py --workload $w $extra --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); r=d['roofline']; e=d['e2e']; c=d['clocks']
print('| %s | %dx%d | %.1f | %d | %d | %.3f | %.2f | %d | %s%s | %.1f |' % (d['config']['name'], d['config']['canvas'][0], d['config']['canvas'][1], d['ms_per_step']*1e3, d['value'], r['achieved'], r['frac'], e['ms_per_step'], e['value'], c['sm_mhz'], ' (power cap)' if 'sw_power_cap' in c['reasons'] else '', d['host_us_per_submit']))"
done
