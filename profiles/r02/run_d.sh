# r02 call D: PNG encoder after the length-limit fix, new bench with legs, A/B of fused vs stand-alone survivor lists
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -12
echo "== resize probe"; timeout 300 python profiles/r02/resize_probe.py
echo "== bench (default: headline + legs + cpu)"; timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_d.json 2> gpurun_out/bench_d.err; tail -3 gpurun_out/bench_d.err; python profiles/r02/summ.py gpurun_out/bench_d.json
for t in 1 2 3; do echo "== headline only, threads $t"; timeout 300 python bench.py --steps 8 --warmup 3 --no-legs --no-cpu --no-check --threads $t > gpurun_out/bench_d_t$t.json 2>/dev/null; python profiles/r02/summ.py gpurun_out/bench_d_t$t.json; done
echo "== stand-alone list kernel (ICB_FUSED_LISTS=0)"; ICB_FUSED_LISTS=0 timeout 300 python bench.py --steps 8 --warmup 3 --no-legs --no-cpu --no-check > gpurun_out/bench_d_unfused.json 2>/dev/null; python profiles/r02/summ.py gpurun_out/bench_d_unfused.json
echo "== --impl reference"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 | cut -c1-400
