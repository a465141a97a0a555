# quick correctness + throughput pass (gpurun, one B200): GPU parity suite, cfg1 bench without the CPU leg, launch list
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_quick.log
tail -4 gpurun_out/pytest_quick.log
timeout 300 python bench.py --no-cpu > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; cat gpurun_out/bench_quick.json
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
ncu --metrics gpu__time_duration.sum --clock-control none -s 120 -c 120 --csv --log-file gpurun_out/launches_quick.csv $CMD > gpurun_out/ncu_launches_quick.log 2>&1
python profiles/launch_summary.py gpurun_out/launches_quick.csv
