B="python bench.py --steps 8 --warmup 3 --no-cpu --no-check"
IMGCOMP_B200_LIB=$PWD/variants_tmp/lib_v1.so $B > gpurun_out/bench_n_v1.json 2>&1; python profiles/summ.py gpurun_out/bench_n_v1.json | tail -1 | cut -c1-40
ICB_PDL=0 $B > gpurun_out/bench_n_nopdl.json 2>&1; python profiles/summ.py gpurun_out/bench_n_nopdl.json | tail -1 | cut -c1-40
ICB_PASS_CPS=3 $B > gpurun_out/bench_n_c3.json 2>&1; python profiles/summ.py gpurun_out/bench_n_c3.json | tail -1 | cut -c1-40
ICB_PASS_CPS=2 $B > gpurun_out/bench_n_c2.json 2>&1; python profiles/summ.py gpurun_out/bench_n_c2.json | tail -1 | cut -c1-40
$B --slots 3 > gpurun_out/bench_n_s3.json 2>&1; python profiles/summ.py gpurun_out/bench_n_s3.json | tail -1 | cut -c1-40
$B --slots 2 > gpurun_out/bench_n_s2.json 2>&1; python profiles/summ.py gpurun_out/bench_n_s2.json | tail -1 | cut -c1-40
