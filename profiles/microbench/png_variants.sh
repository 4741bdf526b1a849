# decode_png wall time of one screenshot / one photo for every library under variants/ (unfilter protocol A/B)
for lib in default $(ls variants/*.so 2>/dev/null); do
  if [ $lib = default ]; then unset STITCH_LIB; else export STITCH_LIB=$PWD/$lib; fi
  echo "== $lib"
  timeout 200 python profiles/microbench/png_decode_time.py shot 0 2>&1 | grep "one file" | tail -1
  timeout 200 python profiles/microbench/png_decode_time.py photo 0 2>&1 | grep "one file" | tail -1
done
