# quick A/B pass: full GPU parity suite, cfg1/cfg2/cfg5 bench without the CPU leg, launch list
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --no-cpu --no-check --steps 8 --warmup 3 > gpurun_out/bench_ab.json 2> gpurun_out/bench_ab.err; python profiles/summ.py gpurun_out/bench_ab.json
timeout 300 python bench.py --no-cpu --no-check --workload cfg2 --batch 32 --steps 6 > gpurun_out/bench_ab_cfg2.json 2>/dev/null; python profiles/summ.py gpurun_out/bench_ab_cfg2.json
timeout 300 python bench.py --no-cpu --no-check --workload cfg5 --steps 10 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('cfg5', d['value'], d['unit'], 'e2e', d['e2e']['value'])"
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
ncu --metrics gpu__time_duration.sum --clock-control none -s 120 -c 120 --csv --log-file gpurun_out/launches_ab.csv $CMD > gpurun_out/ncu_launches_ab.log 2>&1
python profiles/launch_summary.py gpurun_out/launches_ab.csv
