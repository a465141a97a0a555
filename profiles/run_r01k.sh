B="python bench.py --steps 5 --warmup 3 --no-cpu --no-check"
ICB_PASS_CPS=6 $B > gpurun_out/bench_k_c6.json 2>&1
ICB_PASS_CPS=5 $B > gpurun_out/bench_k_c5.json 2>&1
ICB_PASS_CPS=4 $B > gpurun_out/bench_k_c4.json 2>&1
ICB_PASS_CPS=8 $B > gpurun_out/bench_k_c8.json 2>&1
