# profile of pass v7 + prepare + transform (run under gpurun from the repo root)
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 160 -c 100 --csv --log-file gpurun_out/launches_v7.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k "regex:k_entropy_pass" -s 14 -c 2 -o gpurun_out/prof_pass_v7 $CMD > gpurun_out/ncu_pass.log 2>&1
ls -la gpurun_out/ | tail -4
