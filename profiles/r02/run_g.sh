# r02 call G: parity after the PNG / resize / job-start changes, PNG kernel times, N1 A/B, default bench
export ICB_POOL_CACHE=/tmp/icb_pool
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
echo "== resize probe"; timeout 300 python profiles/r02/resize_probe.py
r() { timeout 300 python bench.py --steps 8 --warmup 3 --no-legs --no-cpu --no-check "$@" 2>/dev/null > gpurun_out/g_tmp.json; python profiles/r02/summ.py gpurun_out/g_tmp.json | head -1 | cut -c22-; }
echo "== headline slots 8"; r --slots 8
echo "== N1: bit-length-only pass build (sizes wrong on purpose; timing only)"; IMGCOMP_B200_LIB=$PWD/image_compression_b200/libimgcomp_b200_bits.so r --slots 8
echo "== default bench with legs"; timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_g.json 2> gpurun_out/bench_g.err; tail -3 gpurun_out/bench_g.err; python profiles/r02/summ.py gpurun_out/bench_g.json
CMD="python profiles/r02/png_probe.py"
$CMD > gpurun_out/plain_g.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r02_launches_png.csv $CMD > gpurun_out/ncu_png.log 2>&1
cat gpurun_out/plain_g.log
