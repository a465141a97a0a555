CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_entropy -s 12 -c 2 -o gpurun_out/prof_entropy $CMD > gpurun_out/ncu_entropy.log 2>&1
$CMD > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_transform -s 4 -c 2 -o gpurun_out/prof_transform $CMD > gpurun_out/ncu_transform.log 2>&1
ls -la gpurun_out/
tail -3 gpurun_out/ncu_entropy.log gpurun_out/ncu_transform.log
