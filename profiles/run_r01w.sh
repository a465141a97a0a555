timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_w.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_w.log
tail -3 gpurun_out/pytest_w.log
B="python bench.py --steps 8 --warmup 3 --no-cpu"
$B > gpurun_out/bench_w.json 2>&1; python profiles/summ.py gpurun_out/bench_w.json | tail -1 | cut -c1-200
$B --workload cfg2 --batch 32 > gpurun_out/bench_w_cfg2.json 2>&1; python profiles/summ.py gpurun_out/bench_w_cfg2.json | tail -1 | cut -c1-200
