# grid-size knobs on the aggregate bench value
for cps in 1 2 3; do
  ICB_PASS_CPS=$cps timeout 200 python bench.py --no-cpu > gpurun_out/knob_pass$cps.json 2> gpurun_out/knob_pass$cps.err
done
for cps in 2 3; do
  ICB_XF_CPS=$cps timeout 200 python bench.py --no-cpu > gpurun_out/knob_xf$cps.json 2> gpurun_out/knob_xf$cps.err
done
ICB_PASS_CPS=2 ICB_XF_CPS=2 timeout 200 python bench.py --no-cpu > gpurun_out/knob_both2.json 2> gpurun_out/knob_both2.err
timeout 200 python bench.py --no-cpu --slots 8 > gpurun_out/knob_slots8.json 2> gpurun_out/knob_slots8.err
timeout 200 python bench.py --no-cpu --slots 12 > gpurun_out/knob_slots12.json 2> gpurun_out/knob_slots12.err
timeout 200 python bench.py --no-cpu > gpurun_out/knob_base.json 2> gpurun_out/knob_base.err
grep -H -o '"value": [0-9.]*' gpurun_out/knob_*.json
