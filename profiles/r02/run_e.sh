# r02 call E: where did the resident rate go (12.65 k -> 10.4 k img/s)?  variants x list strategy x D2H method
export ICB_POOL_CACHE=/tmp/icb_pool
timeout 900 python -m pytest tests/test_gpu_queue.py tests/test_gpu_png.py -m gpu -x -q 2>&1 | tail -5
r() { timeout 300 python bench.py --steps 8 --warmup 3 --no-legs --no-cpu --no-check "$@" 2>/dev/null > gpurun_out/e_tmp.json; python profiles/r02/summ.py gpurun_out/e_tmp.json | head -1 | cut -c22-; }
L=$PWD/image_compression_b200/libimgcomp_b200
for lib in "" _ldg _nokclk _t16; do
  for fused in 1 0; do for mc in 0 1; do
    echo "== lib${lib:-_default} fused=$fused fetch_memcpy=$mc"; IMGCOMP_B200_LIB=$L$lib.so ICB_FUSED_LISTS=$fused ICB_FETCH_MEMCPY=$mc r
  done; done
done
echo "== default, slots 8 / 12"; r --slots 8; r --slots 12 --threads 3
echo "== default, pdl off"; ICB_PDL=0 r
