B="python bench.py --steps 8 --warmup 3 --no-cpu --no-check"
ICB_XF_CPS=3 $B > gpurun_out/bench_s_x3.json 2>&1; python profiles/summ.py gpurun_out/bench_s_x3.json | tail -1 | cut -c1-200
ICB_XF_CPS=2 $B > gpurun_out/bench_s_x2.json 2>&1; python profiles/summ.py gpurun_out/bench_s_x2.json | tail -1 | cut -c1-200
ICB_XF_CPS=3 ICB_PASS_CPS=3 $B > gpurun_out/bench_s_x3c3.json 2>&1; python profiles/summ.py gpurun_out/bench_s_x3c3.json | tail -1 | cut -c1-200
ICB_XF_CPS=2 ICB_PASS_CPS=2 $B --slots 8 > gpurun_out/bench_s_x2c2.json 2>&1; python profiles/summ.py gpurun_out/bench_s_x2c2.json | tail -1 | cut -c1-200
