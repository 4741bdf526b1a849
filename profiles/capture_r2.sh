# Round-2 evidence capture (one gpurun call): the launch list of the default bench command and one --set full capture of
# the compose kernel per workload.  Each ncu run follows a plain run of the same command that exited 0.
set -x
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --files-steps 4 --e2e-steps 3"
$B > gpurun_out/plain_launches.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches.csv $B > gpurun_out/ncu_launches.log 2>&1
for w in cfg2 cfg2_yuv cfg3 cfg3_min cfg1; do
  C="python bench.py --workload $w --steps 6 --warmup 3 --no-cpu-baseline --e2e-steps 1 --files-steps 0"
  $C > gpurun_out/plain_$w.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:compose -s 60 -c 1 -o gpurun_out/r2_$w $C > gpurun_out/ncu_$w.log 2>&1
done
ls -la gpurun_out/
