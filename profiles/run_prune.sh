timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pr_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pr_pytest.log
tail -3 gpurun_out/pr_pytest.log
timeout 200 python bench.py --workload cfg5 --steps 10 --no-cpu > gpurun_out/pr_cfg5_on.json 2>&1
ICB_PRUNE_FROM=9 timeout 200 python bench.py --workload cfg5 --steps 10 --no-cpu > gpurun_out/pr_cfg5_off.json 2>&1
ICB_PRUNE_FROM=0.2 timeout 200 python bench.py --no-cpu > gpurun_out/pr_cfg1_on.json 2>&1
timeout 200 python bench.py --no-cpu > gpurun_out/pr_cfg1_off.json 2>&1
grep -H -o '"value": [0-9.]*' gpurun_out/pr_*.json
