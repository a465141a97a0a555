# r02 call M: is the file path bound by the launch rate?  rounds queued 10 / 6 / 3
for r in 10 6 3; do echo "== rounds $r"; ICB_DEC_ROUNDS=$r PROBE_SLOTS=12 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | sed -n 2p; done
