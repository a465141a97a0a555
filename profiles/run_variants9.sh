# is the resident number bound by the host's launch rate?  caller threads x slots
b() { timeout 300 python bench.py --no-cpu --no-check --steps 8 --warmup 3 "$@" 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('   img/s %d e2e %d host_wall %s' % (d['images_per_s'], d['e2e']['images_per_s'], d['host_wall_ms_per_step']))"; }
echo "threads 1 slots 6"; b
echo "threads 2 slots 6"; b --threads 2
echo "threads 3 slots 6"; b --threads 3
echo "threads 2 slots 12"; b --threads 2 --slots 12
echo "threads 4 slots 12"; b --threads 4 --slots 12
echo "threads 1 slots 6 batch 128"; b --batch 128
