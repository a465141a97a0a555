# r02 call N: decode tests after the two-level code table, default bench with the file-input e2e leg
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -4
PROBE_SLOTS=12 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | sed -n 2p
echo "== default bench with legs"; timeout 1200 python bench.py > gpurun_out/bench_n.json 2> gpurun_out/bench_n.err; tail -5 gpurun_out/bench_n.err; python profiles/r02/summ.py gpurun_out/bench_n.json
