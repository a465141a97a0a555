timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_l.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_l.log
tail -5 gpurun_out/pytest_l.log
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_l.json 2> gpurun_out/bench_l.err
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
ncu --metrics gpu__time_duration.sum --clock-control none -s 140 -c 100 --csv --log-file gpurun_out/launches_l.csv $CMD > gpurun_out/ncu_launches_l.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:k_entropy_prepare" -s 3 -c 1 -o gpurun_out/prof_prepare_l $CMD > gpurun_out/ncu_prepare_l.log 2>&1
