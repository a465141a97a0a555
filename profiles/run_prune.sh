timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pr_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pr_pytest.log
tail -3 gpurun_out/pr_pytest.log
timeout 200 python bench.py --workload cfg5 --steps 10 --no-cpu > gpurun_out/pr_cfg5_on.json 2>&1
grep -H -o '"value": [0-9.]*' gpurun_out/pr_*.json
