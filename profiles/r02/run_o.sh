# r02 call O: decode tests
export ICB_POOL_CACHE=/tmp/icb_pool
timeout 900 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -8
