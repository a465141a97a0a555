# r02 call L: what the decode kernels wait for (ncu full), and more jobs in flight
mkdir -p gpurun_out
for sl in 12 16; do echo "== slots $sl"; PROBE_SLOTS=$sl timeout 600 python profiles/r02/decode_probe.py 2>&1 | grep file_fit; done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_dec_write|k_dec_sync" -c 2 -o gpurun_out/r02_decode -f python profiles/r02/decode_ncu.py > gpurun_out/ncu_l.log 2>&1
tail -3 gpurun_out/ncu_l.log
