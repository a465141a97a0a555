# r02 call I: parity after the slot policy + joined EOB/DC code table; A/B of the pass change (same box, alternating),
# the N1 A/B (bit-length-only build of the pass), default bench with legs
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
r() { timeout 300 python bench.py --steps 10 --warmup 3 --no-legs --no-cpu --no-check "$@" 2>/dev/null > gpurun_out/i_tmp.json; python profiles/r02/summ.py gpurun_out/i_tmp.json | head -1 | cut -c22-; }
for k in 1 2; do
  echo "== base"; IMGCOMP_B200_LIB=$PWD/image_compression_b200/libimgcomp_b200_base.so r
  echo "== new";  r
done
echo "== N1: bit-length-only pass (sizes wrong on purpose; timing only)"; IMGCOMP_B200_LIB=$PWD/image_compression_b200/libimgcomp_b200_bits.so r
echo "== new, slots 8"; r --slots 8
echo "== default bench with legs"; timeout 900 python bench.py > gpurun_out/bench_i.json 2> gpurun_out/bench_i.err; tail -3 gpurun_out/bench_i.err; python profiles/r02/summ.py gpurun_out/bench_i.json
