# padded survivor lists (default build) and unconditional window ORs (_orany)
for v in "" _orany; do
  export IMGCOMP_B200_LIB=$PWD/image_compression_b200/libimgcomp_b200$v.so
  echo "=== variant '$v'"
  timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -1
  timeout 300 python bench.py --no-cpu --no-check --steps 8 --warmup 3 > gpurun_out/bench_var$v.json 2> gpurun_out/bench_var$v.err; python profiles/summ.py gpurun_out/bench_var$v.json
  timeout 300 python bench.py --no-cpu --no-check --workload cfg2 --batch 32 --steps 6 > gpurun_out/bench_var${v}_cfg2.json 2>/dev/null; python profiles/summ.py gpurun_out/bench_var${v}_cfg2.json
done
