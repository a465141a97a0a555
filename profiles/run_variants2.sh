# where does the transform's time go: builds without arithmetic / without stores, and grid sizes (ncu launch times only)
t() { ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_transform -s 8 -c 16 --csv --log-file gpurun_out/lv.csv python bench.py --steps 1 --warmup 1 --batch 8 --no-cpu --no-check > /dev/null 2>&1; python profiles/launch_summary.py gpurun_out/lv.csv | tail -1; }
for v in _dbg1 _dbg2 _dbg3 _dbg3s4; do
  export IMGCOMP_B200_LIB=$PWD/image_compression_b200/libimgcomp_b200$v.so
  echo "=== variant '$v'"; t
done
unset IMGCOMP_B200_LIB
for c in 1 2 3 4 6; do echo "=== default build, ICB_XF_CPS=$c"; ICB_XF_CPS=$c t; done
