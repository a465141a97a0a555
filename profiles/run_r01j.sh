timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_j.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_j.log
tail -15 gpurun_out/pytest_j.log
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
ncu --metrics gpu__time_duration.sum --clock-control none -s 140 -c 100 --csv --log-file gpurun_out/launches_j.csv $CMD > gpurun_out/ncu_launches_j.log 2>&1
