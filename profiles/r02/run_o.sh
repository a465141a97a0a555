# r02 call O: blocks cleared as the write pass starts them (no 36 MB clear up front)
timeout 600 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -2
for z in 0 1; do echo "== clear up front: $z"; ICB_DEC_ZERO=$z PROBE_SLOTS=24 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | grep images_per_s | sed -n 3p
echo -n "   reader alone: "; ICB_DEC_ZERO=$z ICB_DEC_ONLY=1 PROBE_SLOTS=24 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | grep images_per_s | sed -n 3p; done
