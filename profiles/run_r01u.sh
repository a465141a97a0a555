timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_u.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_u.log
tail -3 gpurun_out/pytest_u.log
B="python bench.py --steps 8 --warmup 3 --no-cpu"
$B > gpurun_out/bench_u.json 2>&1; python profiles/summ.py gpurun_out/bench_u.json | tail -1 | cut -c1-200
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k "regex:k_entropy_pass" -s 12 -c 3 --csv --log-file gpurun_out/pass_u.csv $CMD > /dev/null 2>&1
tail -9 gpurun_out/pass_u.csv | cut -d, -f5,12- 
