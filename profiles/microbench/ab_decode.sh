# JPEG decoder granularity A/B: file-job throughput (e2e_files) and the decoder parity tests for every library under variants/
for lib in default $(ls variants/*.so 2>/dev/null); do
  if [ $lib = default ]; then unset STITCH_LIB; else export STITCH_LIB=$PWD/$lib; fi
  python bench.py --no-cpu-baseline --steps 20 --warmup 3 --e2e-steps 3 --files-steps 30 2>/dev/null | python -c "
import sys,json
ls=[l for l in sys.stdin if l.startswith('{')]
if not ls: print('$lib FAILED'); sys.exit(0)
f=json.loads(ls[-1])['e2e_files']
print('%-36s png %.2f ms/job (single %.2f)  jpeg %.2f ms/job (single %.2f)' % ('$lib', f['png']['ms_per_step'], f['png']['single_job_latency_ms'], f['jpeg']['ms_per_step'], f['jpeg']['single_job_latency_ms']))"
  timeout 300 python -m pytest tests/test_jpeg_decode.py -x -q -m gpu 2>&1 | tail -1
done
