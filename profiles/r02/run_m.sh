# r02 call M: subsequence 4096 / 8192, carry, slots (32 hardware queues)
timeout 600 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -3
for cfg in "4096 2 24" "4096 3 24" "4096 4 24" "4096 4 32" "8192 2 24" "8192 2 32" "8192 3 32"; do set -- $cfg; echo "== sub $1 carry $2 slots $3"; ICB_DEC_SUB=$1 ICB_DEC_CARRY=$2 PROBE_SLOTS=$3 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | sed -n 2,3p; done
