# entropy tile of 256 vs 128 vs 64 blocks: throughput (cfg1, cfg2) and single-image latency (cfg5)
for v in _t256 "" _t64; do
  export IMGCOMP_B200_LIB=$PWD/image_compression_b200/libimgcomp_b200$v.so
  echo "=== variant '$v'"
  timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py -m gpu -x -q 2>&1 | tail -1
  timeout 300 python bench.py --no-cpu --no-check --steps 8 --warmup 3 > gpurun_out/bench_var$v.json 2> gpurun_out/bench_var$v.err; python profiles/summ.py gpurun_out/bench_var$v.json
  timeout 300 python bench.py --no-cpu --no-check --workload cfg2 --batch 32 --steps 6 > gpurun_out/bench_var${v}_cfg2.json 2>/dev/null; python profiles/summ.py gpurun_out/bench_var${v}_cfg2.json
  timeout 300 python bench.py --no-cpu --no-check --workload cfg5 --steps 10 > gpurun_out/bench_var${v}_cfg5.json 2>/dev/null; python -c "
import json;d=json.loads(open('gpurun_out/bench_var${v}_cfg5.json').read().strip().splitlines()[-1]);print('cfg5', d['metric'], d['value'], d['unit'], 'e2e', d['e2e']['value'])"
done
