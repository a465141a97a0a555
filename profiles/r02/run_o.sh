# r02 call O: bit reader with the next word asked for one refill ahead
timeout 600 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -2
PROBE_SLOTS=24 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | grep images_per_s | sed -n 2,3p
echo "== decode only"; ICB_DEC_ONLY=1 PROBE_SLOTS=24 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | grep images_per_s | sed -n 3p
