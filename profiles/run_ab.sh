# quick A/B pass: parity subset, cfg1 bench without the CPU leg, launch list
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --no-cpu > gpurun_out/bench_ab.json 2> gpurun_out/bench_ab.err; python profiles/summ.py gpurun_out/bench_ab.json
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
ncu --metrics gpu__time_duration.sum --clock-control none -s 120 -c 120 --csv --log-file gpurun_out/launches_ab.csv $CMD > gpurun_out/ncu_launches_ab.log 2>&1
python profiles/launch_summary.py gpurun_out/launches_ab.csv
[ -x profiles/micro/pipe_rates ] && profiles/micro/pipe_rates | tee gpurun_out/pipe_rates.txt
