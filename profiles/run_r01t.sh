timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_t.log
tail -3 gpurun_out/pytest_t.log
B="python bench.py --steps 8 --warmup 3 --no-cpu"
$B > gpurun_out/bench_t.json 2>&1; python profiles/summ.py gpurun_out/bench_t.json | tail -1 | cut -c1-200
