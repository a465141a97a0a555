# r02 call Q: decode chain as a CUDA graph -- parity, host time per submit, rate
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -3
echo "== graph"; ICB_DEC_DEBUG=1 PROBE_SLOTS=24 python profiles/r02/decode_probe2.py 2>&1 | grep -v "icb decode" | sed -n 1,12p
echo "== plain launches"; ICB_DEC_GRAPH=0 ICB_DEC_DEBUG=1 PROBE_SLOTS=24 python profiles/r02/decode_probe2.py 2>&1 | grep -v "icb decode" | sed -n 1,6p
