# r02 call A: sanity of the generalised transform tiling + A/B of 21-MCU tiles
set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
b() { timeout 300 python bench.py --no-cpu --no-check --steps 8 --warmup 3 "$@" 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('   img/s %d e2e %d xf_us %.2f xf_frac %.3f ent_us %.2f' % (d['images_per_s'], d['e2e']['images_per_s'], 1e3*d['roofline_transform']['ms_per_launch'], d['roofline_transform']['frac'], 1e3*d['roofline_entropy']['ms_per_launch']))"; }
echo "== default (16-MCU tiles)"; b; b --workload cfg2 --batch 32
export IMGCOMP_B200_LIB=$PWD/image_compression_b200/libimgcomp_b200_t21.so
echo "== 21-MCU tiles"; timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py -m gpu -x -q 2>&1 | tail -3
b; b --workload cfg2 --batch 32
for c in 3 4 5; do echo "== 21-MCU tiles, ICB_XF_CPS=$c"; ICB_XF_CPS=$c b; done
