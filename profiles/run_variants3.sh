# butterfly column pass vs expanded form; full ncu capture of the entropy pass
t() { ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_transform -s 8 -c 16 --csv --log-file gpurun_out/lv.csv python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check > /dev/null 2>&1; python profiles/launch_summary.py gpurun_out/lv.csv | tail -1; }
for v in "" _bfly; do
  export IMGCOMP_B200_LIB=$PWD/image_compression_b200/libimgcomp_b200$v.so
  echo "=== variant '$v'"; timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1; t
  timeout 300 python bench.py --no-cpu --no-check --steps 6 --warmup 3 > gpurun_out/bench_var$v.json 2> gpurun_out/bench_var$v.err; python profiles/summ.py gpurun_out/bench_var$v.json
done
unset IMGCOMP_B200_LIB
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
ncu --set full --clock-control none --import-source on -k "regex:k_entropy_pass" -s 12 -c 2 -f -o gpurun_out/prof_pass_r2 $CMD > gpurun_out/ncu_pass_r2.log 2>&1
ls -la gpurun_out/*.ncu-rep
