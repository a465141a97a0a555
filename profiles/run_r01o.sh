timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_o.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_o.log
tail -4 gpurun_out/pytest_o.log
B="python bench.py --steps 8 --warmup 3 --no-cpu --no-check"
for v in f cur f cur; do
  if [ $v = cur ]; then L=""; else L="$PWD/variants_tmp/lib_$v.so"; fi
  IMGCOMP_B200_LIB=$L $B > gpurun_out/bench_o_$v.json 2>&1
  python profiles/summ.py gpurun_out/bench_o_$v.json | tail -1 | cut -c1-200
done
