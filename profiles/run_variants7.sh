# throughput knobs: programmatic dependent launch on/off, slots, caller threads
b() { timeout 300 python bench.py --no-cpu --no-check --steps 8 --warmup 3 "$@" 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('   img/s %d e2e %d' % (d['images_per_s'], d['e2e']['images_per_s']))"; }
echo "default"; b
echo "ICB_PDL=0"; ICB_PDL=0 b
echo "ICB_PDL=0 slots 8"; ICB_PDL=0 b --slots 8
echo "ICB_PDL=0 slots 12"; ICB_PDL=0 b --slots 12
echo "slots 4"; b --slots 4
echo "ICB_PDL=0 cfg2"; ICB_PDL=0 b --workload cfg2 --batch 32
echo "cfg2"; b --workload cfg2 --batch 32
