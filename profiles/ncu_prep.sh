CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
timeout 200 $CMD > gpurun_out/plain_prep.log 2>&1 || exit 1
timeout 400 ncu --set full --clock-control none --import-source on -k "regex:k_entropy_prepare" -s 10 -c 1 -f -o gpurun_out/prof_prep_v2 $CMD > gpurun_out/ncu_prep_v2.log 2>&1
IMGCOMP_B200_LIB=$PWD/variants_tmp/lib_v1.so timeout 400 ncu --set full --clock-control none --import-source on -k "regex:k_entropy_prepare" -s 10 -c 1 -f -o gpurun_out/prof_prep_v1 $CMD > gpurun_out/ncu_prep_v1.log 2>&1
ls -la gpurun_out | grep prep
