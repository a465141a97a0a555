# r02 call U: default bench with the file variants of the cfg2 and cfg4 legs
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
timeout 1200 python bench.py > gpurun_out/bench_u.json 2> gpurun_out/bench_u.err; tail -3 gpurun_out/bench_u.err; python profiles/r02/summ.py gpurun_out/bench_u.json > gpurun_out/bench_u.txt 2>&1; cat gpurun_out/bench_u.txt
