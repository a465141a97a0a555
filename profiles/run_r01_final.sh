# round-1 final measurement set (gpurun, one B200)
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_final.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_final.log
tail -3 gpurun_out/pytest_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; tail -1 gpurun_out/smoke_final.log
timeout 400 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_final_ref.json 2>&1
timeout 400 python bench.py --workload cfg2 --batch 32 > gpurun_out/bench_final_cfg2.json 2>&1
timeout 300 python bench.py --workload png4k --steps 5 > gpurun_out/bench_final_png4k.json 2>&1
timeout 300 python bench.py --workload png8k --steps 5 --batch 32 > gpurun_out/bench_final_png8k.json 2>&1
timeout 300 python bench.py --workload cfg4 --steps 5 > gpurun_out/bench_final_cfg4.json 2>&1
timeout 300 python bench.py --workload cfg5 --steps 10 > gpurun_out/bench_final_cfg5.json 2>&1
nproc
for f in gpurun_out/bench_final*.json; do echo $f; tail -1 $f | cut -c1-400; done
