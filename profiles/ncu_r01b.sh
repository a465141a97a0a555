# round-1 profile pass of the current kernels (run under gpurun from the repo root)
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 256 -c 160 --csv --log-file gpurun_out/launches_v3.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_entropy_symbols -s 14 -c 2 -o gpurun_out/prof_symbols_v3 $CMD > gpurun_out/ncu_symbols.log 2>&1
$CMD > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_transform -s 4 -c 2 -o gpurun_out/prof_transform_v3 $CMD > gpurun_out/ncu_transform.log 2>&1
$CMD > gpurun_out/plain4.log 2>&1 && ncu --set full --clock-control none --import-source on -k "regex:k_entropy_count|k_entropy_emit" -s 14 -c 3 -o gpurun_out/prof_count_v3 $CMD > gpurun_out/ncu_count.log 2>&1
ls -la gpurun_out/
tail -3 gpurun_out/ncu_symbols.log gpurun_out/ncu_transform.log gpurun_out/ncu_count.log
