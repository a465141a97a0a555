# r02 call P (2 GPUs, last): file jobs through one handle over both GPUs
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -4
