# r02 call O: the reader alone (no fit behind it), slots 24 / 48; then the whole path again
for sl in 24 48; do echo "== decode only, slots $sl"; ICB_DEC_ONLY=1 PROBE_SLOTS=$sl timeout 600 python profiles/r02/decode_probe2.py 2>&1 | grep images_per_s | sed -n 3p; done
echo "== whole path, slots 24"; PROBE_SLOTS=24 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | grep images_per_s | sed -n 3p
