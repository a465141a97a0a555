# concurrent timeline of the last fit of every slot (ICB_TIMING build, GPU global timer), 6 slots sharing one GPU
export IMGCOMP_B200_LIB=$PWD/image_compression_b200/libimgcomp_b200_timing.so
ICB_DUMP_TIMELINE=1 timeout 300 python bench.py --no-cpu --no-check --steps 4 --warmup 2 > gpurun_out/bench_timeline.json 2> gpurun_out/timeline.txt
grep -E "^slot" gpurun_out/timeline.txt | tail -12
python profiles/summ.py gpurun_out/bench_timeline.json
