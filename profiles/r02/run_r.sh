# r02 call R: whole GPU suite, smoke, default bench (graph decode, 24 file slots), reference arm
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 1200 python bench.py > gpurun_out/bench_r.json 2> gpurun_out/bench_r.err; tail -3 gpurun_out/bench_r.err; python profiles/r02/summ.py gpurun_out/bench_r.json > gpurun_out/bench_r.txt 2>&1; cat gpurun_out/bench_r.txt
