# prune/half-list knob on the q=0.25 workloads
ICB_PRUNE_FROM=0.2 timeout 200 python bench.py --no-cpu > gpurun_out/knob_cfg1_prune.json 2> gpurun_out/knob_cfg1_prune.err
timeout 200 python bench.py --no-cpu > gpurun_out/knob_cfg1_base.json 2> gpurun_out/knob_cfg1_base.err
ICB_PRUNE_FROM=0.2 timeout 200 python bench.py --no-cpu --workload cfg2 --batch 32 > gpurun_out/knob_cfg2_prune.json 2> gpurun_out/knob_cfg2_prune.err
timeout 200 python bench.py --no-cpu --workload cfg2 --batch 32 > gpurun_out/knob_cfg2_base.json 2> gpurun_out/knob_cfg2_base.err
grep -H -o '"value": [0-9.]*' gpurun_out/knob_cfg*.json
grep -h -o '"parity_check": "[^"]*"' gpurun_out/knob_cfg2_base.json
