B="python bench.py --steps 5 --warmup 3 --no-cpu --no-check"
$B --threads 2 > gpurun_out/bench_g_t2.json 2>&1
$B --threads 3 > gpurun_out/bench_g_t3.json 2>&1
$B --threads 3 --slots 12 --batch 96 > gpurun_out/bench_g_t3s12.json 2>&1
$B --threads 6 --slots 12 --batch 96 > gpurun_out/bench_g_t6s12.json 2>&1
ICB_PASS_CPS=2 $B --threads 3 --slots 12 --batch 96 > gpurun_out/bench_g_t3s12c2.json 2>&1
