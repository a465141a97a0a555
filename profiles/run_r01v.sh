timeout 600 python -m pytest tests/test_gpu_edges.py -m gpu -x -q > gpurun_out/pytest_v.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_v.log
tail -15 gpurun_out/pytest_v.log
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_edges.py -m gpu -x -q -k "padded or tiny or bgr or alpha" > gpurun_out/sanitizer_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/sanitizer_memcheck.log
tail -6 gpurun_out/sanitizer_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "test_fit_cached_params or test_encode_stuffing_stress" > gpurun_out/sanitizer_racecheck.log 2>&1; echo "racecheck rc=$?" >> gpurun_out/sanitizer_racecheck.log
tail -12 gpurun_out/sanitizer_racecheck.log
