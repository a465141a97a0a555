CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_entropy -s 12 -c 2 -o gpurun_out/prof_entropy2 $CMD > gpurun_out/ncu_entropy2.log 2>&1
ls -la gpurun_out/ | tail -5
