# r02 call H (re-entry): the no-alloc test alone then the whole GPU suite, smoke, default bench with legs,
# launch list of a short bench, full ncu of the fit kernels and the resize kernel
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
echo "== no-alloc test alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5
echo "== whole GPU suite"; timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
echo "== default bench with legs"; timeout 1200 python bench.py > gpurun_out/bench_h.json 2> gpurun_out/bench_h.err; tail -3 gpurun_out/bench_h.err; python profiles/r02/summ.py gpurun_out/bench_h.json
echo "== reference arm"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_h_ref.json 2>> gpurun_out/bench_h.err; cat gpurun_out/bench_h_ref.json | cut -c1-400
echo "== resize probe"; timeout 300 python profiles/r02/resize_probe.py
CMD="python bench.py --steps 1 --warmup 1 --batch 16 --no-legs --no-cpu --no-check"
$CMD > gpurun_out/plain_h.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 200 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_entropy_pass|k_transform|k_entropy_emit|k_job_begin" -s 60 -c 12 -o gpurun_out/r02_fit -f $CMD > gpurun_out/ncu_fit.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_resize -s 30 -c 1 -o gpurun_out/r02_resize_rgba4k -f python profiles/r02/resize_probe.py > gpurun_out/ncu_resize.log 2>&1
ls -la gpurun_out | head -30
