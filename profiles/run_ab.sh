# A/B of the in-tree library against variants_tmp/*.so on the aggregate bench value
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/ab_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ab_pytest.log
tail -3 gpurun_out/ab_pytest.log
for rep in 1 2; do
  timeout 200 python bench.py > gpurun_out/ab_new_$rep.json 2> gpurun_out/ab_new_$rep.err
  for v in variants_tmp/*.so; do
    n=$(basename $v .so)
    IMGCOMP_B200_LIB=$PWD/$v timeout 200 python bench.py > gpurun_out/ab_${n}_$rep.json 2> gpurun_out/ab_${n}_$rep.err
  done
done
timeout 200 python bench.py --workload cfg2 --batch 32 > gpurun_out/ab_new_cfg2.json 2>&1
timeout 200 python bench.py --workload cfg5 --steps 10 > gpurun_out/ab_new_cfg5.json 2>&1
grep -h -o '"value": [0-9.]*' gpurun_out/ab_*.json | head -20
