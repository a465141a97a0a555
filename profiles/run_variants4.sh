# entropy tile of 128 blocks vs 64; window zeroing by 16-byte stores (default build already has it)
for v in "" _t128; do
  export IMGCOMP_B200_LIB=$PWD/image_compression_b200/libimgcomp_b200$v.so
  echo "=== variant '$v'"
  timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -1
  timeout 300 python bench.py --no-cpu --steps 8 --warmup 3 > gpurun_out/bench_var$v.json 2> gpurun_out/bench_var$v.err; python profiles/summ.py gpurun_out/bench_var$v.json
  ncu --metrics gpu__time_duration.sum --clock-control none -s 120 -c 120 --csv --log-file gpurun_out/launches_var$v.csv python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check > /dev/null 2>&1
  python profiles/launch_summary.py gpurun_out/launches_var$v.csv
  timeout 300 python bench.py --no-cpu --no-check --steps 8 --warmup 3 --slots 8 > gpurun_out/bench_var${v}_s8.json 2>/dev/null; python profiles/summ.py gpurun_out/bench_var${v}_s8.json
  timeout 300 python bench.py --no-cpu --no-check --workload cfg2 --batch 32 --steps 6 > gpurun_out/bench_var${v}_cfg2.json 2>/dev/null; python profiles/summ.py gpurun_out/bench_var${v}_cfg2.json
done
