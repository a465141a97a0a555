timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pr_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pr_pytest.log
tail -3 gpurun_out/pr_pytest.log
timeout 200 python bench.py --workload cfg5 --steps 10 --no-cpu > gpurun_out/pr_cfg5.json 2>&1
timeout 200 python bench.py --no-cpu > gpurun_out/pr_cfg1.json 2>&1
timeout 200 python bench.py --no-cpu --workload cfg2 --batch 32 > gpurun_out/pr_cfg2.json 2>&1
ICB_PRUNE=1 timeout 200 python -m pytest tests -m gpu -x -q -k "fuzz or golden or fit" > gpurun_out/pr_pytest_force1.log 2>&1; tail -1 gpurun_out/pr_pytest_force1.log
ICB_PRUNE=0 timeout 200 python -m pytest tests -m gpu -x -q -k "fuzz or golden or fit" > gpurun_out/pr_pytest_force0.log 2>&1; tail -1 gpurun_out/pr_pytest_force0.log
grep -H -o '"value": [0-9.]*' gpurun_out/pr_cfg*.json
