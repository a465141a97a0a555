B="python bench.py --steps 8 --warmup 3 --no-cpu --no-check"
for v in f j cur f j cur; do
  if [ $v = cur ]; then L=""; else L="$PWD/variants_tmp/lib_$v.so"; fi
  IMGCOMP_B200_LIB=$L $B > gpurun_out/bench_m_$v.json 2>&1
  python profiles/summ.py gpurun_out/bench_m_$v.json | tail -1 | cut -c1-40
done
