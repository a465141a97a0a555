CMD="python bench.py --workload cfg5 --steps 1 --warmup 1 --no-cpu --no-check"
timeout 200 $CMD > gpurun_out/plain_cfg5.log 2>&1 || exit 1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_cfg5.csv $CMD > gpurun_out/ncu_launches_cfg5.log 2>&1
tail -2 gpurun_out/ncu_launches_cfg5.log
