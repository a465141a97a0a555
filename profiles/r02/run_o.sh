# r02 call O: decode tests, rate
export ICB_POOL_CACHE=/tmp/icb_pool
timeout 900 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -4
PROBE_SLOTS=24 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | sed -n 2,4p
