# runtime tile size: full GPU suite, cfg1 / cfg2 / cfg5
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --no-cpu --steps 8 --warmup 3 > gpurun_out/bench_rt.json 2> gpurun_out/bench_rt.err; python profiles/summ.py gpurun_out/bench_rt.json
timeout 300 python bench.py --no-cpu --no-check --workload cfg2 --batch 32 --steps 6 > gpurun_out/bench_rt_cfg2.json 2>/dev/null; python profiles/summ.py gpurun_out/bench_rt_cfg2.json
timeout 300 python bench.py --no-cpu --no-check --workload cfg5 --steps 10 > gpurun_out/bench_rt_cfg5.json 2>/dev/null; python -c "
import json;d=json.loads(open('gpurun_out/bench_rt_cfg5.json').read().strip().splitlines()[-1]);print('cfg5', d['metric'], d['value'], d['unit'], 'e2e', d['e2e']['value'])"
for t in 128 256; do ICB_ENT_TILE=$t timeout 300 python bench.py --no-cpu --no-check --workload cfg5 --steps 10 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('cfg5 tile $t', d['value'], d['unit'], 'e2e', d['e2e']['value'])"; done
