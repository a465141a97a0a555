# full ncu capture of the transform kernel (after a clean run of the same command)
CMD="python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check"
$CMD > gpurun_out/plain_xf.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k "regex:k_transform" -s 10 -c 1 -f -o gpurun_out/prof_transform_xf $CMD > gpurun_out/ncu_transform_xf.log 2>&1
ls -la gpurun_out | grep xf
