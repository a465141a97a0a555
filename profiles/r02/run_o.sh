# r02 call O (last): the fall-back forms of the reader pass the same tests -- plain launches instead of the graph, unbuffered write pass
ICB_DEC_GRAPH=0 timeout 300 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -2
ICB_DEC_WBUF=0 timeout 300 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -2
