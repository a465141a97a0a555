timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_q.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_q.log
tail -3 gpurun_out/pytest_q.log
B="python bench.py --steps 8 --warmup 3 --no-cpu"
$B > gpurun_out/bench_q.json 2>&1; python profiles/summ.py gpurun_out/bench_q.json | tail -1 | cut -c1-200
ICB_PASS_CPS=5 $B > gpurun_out/bench_q_c5.json 2>&1; python profiles/summ.py gpurun_out/bench_q_c5.json | tail -1 | cut -c1-200
ICB_PASS_CPS=3 $B > gpurun_out/bench_q_c3.json 2>&1; python profiles/summ.py gpurun_out/bench_q_c3.json | tail -1 | cut -c1-200
$B --slots 8 > gpurun_out/bench_q_s8.json 2>&1; python profiles/summ.py gpurun_out/bench_q_s8.json | tail -1 | cut -c1-200
$B --slots 4 > gpurun_out/bench_q_s4.json 2>&1; python profiles/summ.py gpurun_out/bench_q_s4.json | tail -1 | cut -c1-200
