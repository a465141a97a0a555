# r02 call O (last): decode tests and smoke after the per-GPU attribute change
timeout 300 python -m pytest tests/test_gpu_decode.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -1
