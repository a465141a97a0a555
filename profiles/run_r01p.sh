B="python bench.py --steps 8 --warmup 3 --no-cpu"
for v in w1 w2 w12 cur; do
  if [ $v = cur ]; then L=""; else L="$PWD/variants_tmp/lib_$v.so"; fi
  IMGCOMP_B200_LIB=$L $B > gpurun_out/bench_p_$v.json 2>&1
  python profiles/summ.py gpurun_out/bench_p_$v.json | tail -1 | cut -c1-200
done
