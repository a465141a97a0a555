# A/B of transform-kernel builds (IMGCOMP_B200_LIB picks the build): parity subset on each, resident bench, ncu launch times
for v in "" _base _c5 _c6 _c7 _c8s1 _c6s1; do
  lib=image_compression_b200/libimgcomp_b200$v.so
  [ -f $lib ] || continue
  export IMGCOMP_B200_LIB=$PWD/$lib
  echo "=== variant '$v'"
  timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1
  timeout 300 python bench.py --no-cpu --no-check --steps 6 --warmup 3 > gpurun_out/bench_var$v.json 2> gpurun_out/bench_var$v.err; python profiles/summ.py gpurun_out/bench_var$v.json
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_transform -s 8 -c 16 --csv --log-file gpurun_out/launches_var$v.csv python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check > /dev/null 2>&1
  python profiles/launch_summary.py gpurun_out/launches_var$v.csv | tail -1
done
