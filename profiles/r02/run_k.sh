# r02 call K: launch list of the GPU JPEG reader (ncu, serialised, cold caches)
mkdir -p gpurun_out
python profiles/r02/decode_ncu.py > gpurun_out/plain_k.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_decode.csv python profiles/r02/decode_ncu.py > gpurun_out/ncu_k.log 2>&1
cat gpurun_out/plain_k.log
python profiles/launch_summary.py gpurun_out/r02_launches_decode.csv
