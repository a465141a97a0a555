# refresh of the tracked bench lines after the last change (default bench with its CPU leg, cfg5)
timeout 400 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; python profiles/summ.py gpurun_out/bench_final.json
timeout 300 python bench.py --workload cfg5 --steps 10 > gpurun_out/bench_final_cfg5.json 2>&1; tail -1 gpurun_out/bench_final_cfg5.json | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
