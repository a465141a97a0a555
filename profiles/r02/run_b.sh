# r02 call B: fused survivor lists (21-MCU units), staged resize kernel, queue API -- parity first, then numbers
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
b() { timeout 300 python bench.py --no-cpu --no-check --steps 8 --warmup 3 "$@" 2>gpurun_out/b_err.log | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);r=d.get('roofline_transform');e=d.get('roofline_entropy')
print('   value %.1f %s img/s %s e2e %s | xf_us %s ent_us %s | roof %s' % (d['value'], d['unit'], d.get('images_per_s'), d['e2e'].get('images_per_s', d['e2e']['value']), r and round(1e3*r['ms_per_launch'],2), e and round(1e3*e['ms_per_launch'],2), d['roofline'] and round(d['roofline']['frac'],3)))" || tail -5 gpurun_out/b_err.log; }
echo "== cfg1"; b
echo "== cfg2"; b --workload cfg2 --batch 32
echo "== cfg5"; b --workload cfg5
echo "== png4k"; b --workload png4k
echo "== png8k"; b --workload png8k
echo "== cfg4"; b --workload cfg4 --batch 40
