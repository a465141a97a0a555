timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_i.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_i.log
tail -15 gpurun_out/pytest_i.log
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_i.json 2> gpurun_out/bench_i.err
python bench.py --workload cfg5 --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_i_cfg5.json 2>&1
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
ncu --metrics gpu__time_duration.sum --clock-control none -s 140 -c 100 --csv --log-file gpurun_out/launches_i.csv $CMD > gpurun_out/ncu_launches_i.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:k_entropy_pass" -s 14 -c 2 -o gpurun_out/prof_pass_i $CMD > gpurun_out/ncu_pass_i.log 2>&1
