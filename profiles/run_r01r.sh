timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_r.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_r.log
tail -4 gpurun_out/pytest_r.log
B="python bench.py --steps 8 --warmup 3 --no-cpu"
$B > gpurun_out/bench_r.json 2>&1; python profiles/summ.py gpurun_out/bench_r.json | tail -1 | cut -c1-200
ICB_PASS_CPS=3 $B > gpurun_out/bench_r_c3.json 2>&1; python profiles/summ.py gpurun_out/bench_r_c3.json | tail -1 | cut -c1-200
