timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_x.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_x.log
tail -5 gpurun_out/pytest_x.log
B="timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu"
$B > gpurun_out/bench_x.json 2>&1; python profiles/summ.py gpurun_out/bench_x.json | tail -1 | cut -c1-200
$B --workload cfg2 --batch 32 > gpurun_out/bench_x_cfg2.json 2>&1; python profiles/summ.py gpurun_out/bench_x_cfg2.json | tail -1 | cut -c1-200
timeout 200 python bench.py --workload cfg5 --steps 10 --no-cpu > gpurun_out/bench_x_cfg5.json 2>&1; tail -1 gpurun_out/bench_x_cfg5.json | cut -c1-120
