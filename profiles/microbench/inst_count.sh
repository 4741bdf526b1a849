# warp instructions / duration / issue utilisation of the compose kernel for each library variant (ncu, one metric pass)
for w in "$@"; do
  for lib in default $(ls variants/*.so 2>/dev/null); do
    if [ $lib = default ]; then unset STITCH_LIB; else export STITCH_LIB=$PWD/$lib; fi
    C="python bench.py --workload $w --steps 6 --warmup 3 --no-cpu-baseline --e2e-steps 1"
    $C > gpurun_out/plain.log 2>&1 && ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none -k regex:compose_kernel -s 60 -c 1 --csv $C 2>/dev/null | grep compose_kernel | awk -F'","' -v t="$w $lib" '{gsub(/"/,"",$NF); printf "%s %s=%s\n", t, $(NF-2), $NF}'
  done
done
