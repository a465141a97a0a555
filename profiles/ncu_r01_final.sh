# round-1 final profile set (run under gpurun from the repo root; each capture follows a clean run of the same command)
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
$CMD > gpurun_out/plain_final.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 120 -c 120 --csv --log-file gpurun_out/launches_final.csv $CMD > gpurun_out/ncu_launches_final.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:k_entropy_pass" -s 12 -c 2 -f -o gpurun_out/prof_pass_final $CMD > gpurun_out/ncu_pass_final.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:k_transform" -s 10 -c 1 -f -o gpurun_out/prof_transform_final $CMD > gpurun_out/ncu_transform_final.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:k_entropy_prepare" -s 10 -c 1 -f -o gpurun_out/prof_prepare_final $CMD > gpurun_out/ncu_prepare_final.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:k_entropy_emit" -s 10 -c 1 -f -o gpurun_out/prof_emit_final $CMD > gpurun_out/ncu_emit_final.log 2>&1
ncu --set full --clock-control none -k "regex:k_resize" -c 2 -f -o gpurun_out/prof_resize_final python bench.py --workload png8k --steps 1 --warmup 1 --batch 4 > gpurun_out/ncu_resize_final.log 2>&1
ls -la gpurun_out | grep final
