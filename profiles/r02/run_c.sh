# r02 call C: parity of everything new (lists in the transform, queue API, PNG encoder), numbers, ncu of transform + resize
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
b() { timeout 300 python bench.py --no-cpu --no-check --steps 8 --warmup 3 "$@" 2>gpurun_out/b_err.log | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);r=d.get('roofline_transform');e=d.get('roofline_entropy')
print('   value %.1f %s img/s %s e2e %s | xf_us %s ent_us %s | roof %s' % (d['value'], d['unit'], d.get('images_per_s'), d['e2e'].get('images_per_s', d['e2e']['value']), r and round(1e3*r['ms_per_launch'],2), e and round(1e3*e['ms_per_launch'],2), d['roofline'] and round(d['roofline']['frac'],3)))" || tail -5 gpurun_out/b_err.log; }
echo "== cfg1"; b
echo "== cfg1 threads 2"; b --threads 2
echo "== cfg2"; b --workload cfg2 --batch 32
echo "== cfg5"; b --workload cfg5
echo "== cfg4"; b --workload cfg4 --batch 40
echo "== resize probe"; timeout 300 python profiles/r02/resize_probe.py
CMD="python profiles/r02/resize_probe.py"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_resize -s 40 -c 2 -o gpurun_out/r02_resize -f $CMD > gpurun_out/ncu_resize.log 2>&1
CMD2="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_transform -s 20 -c 2 -o gpurun_out/r02_xf -f $CMD2 > gpurun_out/ncu_xf.log 2>&1
ls -la gpurun_out
