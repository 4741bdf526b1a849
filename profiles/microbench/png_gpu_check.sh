# PNG decoder check on the GPU box: statistics, parity tests, job timings, launch list of one photo decode
timeout 300 python profiles/microbench/png_debug.py 2>&1 | tail -8
timeout 300 python -m pytest tests/test_png_decode.py -x -q -m gpu 2>&1 | tail -5
timeout 300 python profiles/microbench/png_decode_time.py photo 9 2>&1 | tail -7
timeout 200 python profiles/microbench/png_decode_time.py shot 12 2>&1 | tail -7
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 50 --csv --log-file gpurun_out/png_launches.csv python profiles/microbench/png_decode_time.py ${1:-photo} 1 > gpurun_out/png_ncu.log 2>&1
