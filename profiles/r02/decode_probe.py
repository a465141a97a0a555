"""Device time of the GPU JPEG reader on a 12 MP file (Pillow q=92 4:2:0, the bench's input-staging file) and of
decode + fit from the file, against the pixel path: CUDA-event time over the slot streams (icb_mark)."""
import io, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import image_compression_b200 as pkg
from image_compression_b200 import synth
from PIL import Image

w, h = 4000, 3000
imgs = [synth.natural_image(3000 + i, w, h) for i in range(4)]
files = []
for im in imgs:
    bio = io.BytesIO(); Image.fromarray(im).save(bio, "JPEG", quality=92, subsampling=2); files.append(np.frombuffer(bio.getvalue(), np.uint8))
SLOTS = int(os.environ.get("PROBE_SLOTS", "6"))
with pkg.Engine(devices=[0], slots_per_gpu=SLOTS, max_w=w, max_h=h) as eng:
    ref = np.asarray(Image.open(io.BytesIO(files[0].tobytes())).convert("RGB"))
    got = eng.jpg_decode(files[0])
    print(json.dumps({"file_bytes": int(files[0].size), "decode_equals_libjpeg_turbo": bool(np.array_equal(got, ref))}), flush=True)
    pins = []
    for f in files:
        p = eng.alloc_pinned((f.size,)); p[:] = f; pins.append(p)
    outs = [eng.alloc_pinned((1 << 20,)) for _ in range(2 * SLOTS)]
    def run(n, inflight):
        pend = []
        for i in range(n):
            if len(pend) == inflight:
                eng.wait(pend.pop(0))
            pend.append(eng.submit_jpg_file_fit(pins[i % 4], 0.25, 1 << 20, outs[i % (2 * SLOTS)]))
        for t in pend:
            eng.wait(t)
    run(2 * SLOTS, SLOTS)
    for inflight in (1, SLOTS):
        n = 8 if inflight == 1 else 16 * SLOTS
        eng.mark(0, 0); t0 = time.perf_counter(); run(n, inflight); eng.mark(0, 1)
        ms = eng.elapsed_ms(0); wall = (time.perf_counter() - t0) * 1e3
        print(json.dumps({"file_fit_in_flight": inflight, "images": n, "device_ms_per_image": round(ms / n, 4),
                          "images_per_s": round(n / ms * 1e3, 1), "host_wall_ms": round(wall, 1)}), flush=True)
