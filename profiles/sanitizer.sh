# compute-sanitizer over the kernels with the most to prove (ONE tool per gpurun call: B200_PROFILING.md):
#   bash profiles/sanitizer.sh memcheck|racecheck|initcheck|synccheck > gpurun_out/sanitizer_<tool>.txt
# memcheck : exact-size device allocations (tests/test_gpu_exact_alloc.py), the streaming unit's TMA ring, codecs
# racecheck: the per-warp mbarrier ring (async-proxy writes vs LDS reads), the JPEG / PNG shared-memory bit windows
# initcheck: ring rows read as padding, staging pads
tool=${1:-memcheck}
tests="tests/test_gpu_exact_alloc.py tests/test_gpu_stream.py::test_mixed_ratio_kernel_shapes tests/test_gpu_parity.py::test_randomized_jobs tests/test_async_files.py::test_submitted_bitmap_job_with_file_output_and_errors_at_wait tests/test_jpeg_decode.py::test_job_with_jpeg_files_equals_job_with_decoded_bitmaps"
[ "$tool" = racecheck ] && tests="tests/test_gpu_exact_alloc.py::test_streaming_units_on_tight_device_bitmaps tests/test_async_files.py::test_submitted_bitmap_job_with_file_output_and_errors_at_wait tests/test_jpeg_decode.py::test_job_with_jpeg_files_equals_job_with_decoded_bitmaps"
extra=""
[ "$tool" = memcheck ] && extra="--leak-check no"
[ "$tool" = initcheck ] && extra="--track-unused-memory no"
timeout 1500 compute-sanitizer --tool $tool $extra --print-limit 30 --error-exitcode 0 python -m pytest $tests -x -q -m gpu -p no:cacheprovider 2>&1 | grep -v "^$" | tail -60
