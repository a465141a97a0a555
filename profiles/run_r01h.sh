timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_h.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_h.log
tail -15 gpurun_out/pytest_h.log
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_h.json 2> gpurun_out/bench_h.err
python bench.py --workload cfg2 --steps 3 --warmup 3 --batch 32 --no-cpu > gpurun_out/bench_h_cfg2.json 2>&1
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
ncu --set full --clock-control none --import-source on -k "regex:k_transform" -s 12 -c 2 -o gpurun_out/prof_transform_h $CMD > gpurun_out/ncu_transform_h.log 2>&1
