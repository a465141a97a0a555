# r02 call O: scan / unstuff on aligned 16-byte groups, output through shared memory
timeout 600 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -6
PROBE_SLOTS=24 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | grep images_per_s | sed -n 3p
echo -n "   reader alone: "; ICB_DEC_ONLY=1 PROBE_SLOTS=24 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | grep images_per_s | sed -n 3p
