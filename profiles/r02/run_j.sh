# r02 call J: the GPU JPEG reader -- parity tests, rounds needed, timing, launch list
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
ICB_DEC_DEBUG=1 timeout 900 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -30
echo "== decode probe"; ICB_DEC_DEBUG=1 timeout 600 python profiles/r02/decode_probe.py 2>&1 | sort | uniq -c | sort -rn | head -12
for sub in 1024 4096; do echo "== sub $sub"; ICB_DEC_SUB=$sub timeout 600 python profiles/r02/decode_probe.py 2>&1 | grep file_fit; done
python profiles/r02/decode_ncu.py > gpurun_out/plain_k.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_decode.csv python profiles/r02/decode_ncu.py > gpurun_out/ncu_k.log 2>&1
python profiles/launch_summary.py gpurun_out/r02_launches_decode.csv
