# r02 call F: parity, persistent resize kernel, slots x PDL policy, ncu of resize / pass / transform, launch list
export ICB_POOL_CACHE=/tmp/icb_pool
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
echo "== resize probe"; timeout 300 python profiles/r02/resize_probe.py
r() { timeout 300 python bench.py --steps 8 --warmup 3 --no-legs --no-cpu --no-check "$@" 2>/dev/null > gpurun_out/f_tmp.json; python profiles/r02/summ.py gpurun_out/f_tmp.json | head -1 | cut -c22-; }
for slots in 6 8 12; do for pdl in auto 0 1; do
  echo "== slots $slots pdl $pdl"; if [ $pdl = auto ]; then r --slots $slots; else ICB_PDL=$pdl r --slots $slots; fi
done; done
echo "== slots 12 threads 3"; r --slots 12 --threads 3
echo "== slots 12 threads 4"; r --slots 12 --threads 4
echo "== cfg2 slots 6 / 12"; r --workload cfg2 --batch 32; r --workload cfg2 --batch 32 --slots 12
CMD="python bench.py --steps 1 --warmup 1 --batch 16 --no-legs --no-cpu --no-check"
$CMD > gpurun_out/plain_f.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 200 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_entropy_pass|k_transform|k_entropy_emit|k_job_begin" -s 60 -c 12 -o gpurun_out/r02_fit -f $CMD > gpurun_out/ncu_fit.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_resize -s 30 -c 1 -o gpurun_out/r02_resize_rgba4k -f python profiles/r02/resize_probe.py > gpurun_out/ncu_resize.log 2>&1
ls -la gpurun_out | head -30
