# cycle f: prologue from global, begin merged into prepare, PDL chain; knobs A/B
set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_f.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_f.log
tail -5 gpurun_out/pytest_f.log
B="python bench.py --steps 5 --warmup 3 --no-cpu --no-check"
python bench.py > gpurun_out/bench_f.json 2> gpurun_out/bench_f.err
ICB_PDL=0 $B > gpurun_out/bench_f_nopdl.json 2>&1
ICB_PASS_CPS=2 $B > gpurun_out/bench_f_cps2.json 2>&1
ICB_PASS_CPS=3 $B > gpurun_out/bench_f_cps3.json 2>&1
ICB_PASS_CPS=2 $B --slots 12 > gpurun_out/bench_f_cps2_s12.json 2>&1
python bench.py --workload cfg2 --steps 3 --warmup 3 --batch 32 --no-cpu > gpurun_out/bench_f_cfg2.json 2>&1
python bench.py --workload png4k --steps 3 --warmup 3 > gpurun_out/bench_f_png4k.json 2>&1
python bench.py --workload png8k --steps 3 --warmup 3 --batch 32 > gpurun_out/bench_f_png8k.json 2>&1
python bench.py --workload cfg5 --steps 5 --warmup 3 > gpurun_out/bench_f_cfg5.json 2>&1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_f_ref.json 2>&1
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
ncu --metrics gpu__time_duration.sum --clock-control none -s 140 -c 100 --csv --log-file gpurun_out/launches_f.csv $CMD > gpurun_out/ncu_launches_f.log 2>&1
ls -la gpurun_out | tail -15
