"""File items through the native executor (icb_run_list): threads x jobs in flight, 12 MP q92 files."""
import io, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import image_compression_b200 as pkg
from image_compression_b200 import synth
from PIL import Image

w, h = 4000, 3000
files = []
for i in range(4):
    bio = io.BytesIO(); Image.fromarray(synth.natural_image(3000 + i, w, h)).save(bio, "JPEG", quality=92, subsampling=2)
    files.append(np.frombuffer(bio.getvalue(), np.uint8))
SLOTS = int(os.environ.get("PROBE_SLOTS", "12"))
with pkg.Engine(devices=[0], slots_per_gpu=SLOTS, max_w=w, max_h=h) as eng:
    pins = []
    for f in files:
        p = eng.alloc_pinned((f.size,)); p[:] = f; pins.append(p)
    n = 256
    outs = [eng.alloc_pinned((1 << 20,)) for _ in range(32)]
    items = eng.make_items([dict(kind="jpg_file_fit", file=pins[i % 4], init_q=0.25, target=1 << 20, out=outs[i % 32]) for i in range(n)])
    for threads in (1, 2, 3, 4, 6):
        inflight = max(1, SLOTS // threads)
        eng.run_list(items, threads=threads, inflight=inflight)
        eng.mark(0, 0); t0 = time.perf_counter(); eng.run_list(items, threads=threads, inflight=inflight); eng.mark(0, 1)
        ms = eng.elapsed_ms(0); wall = (time.perf_counter() - t0) * 1e3
        bad = sum(1 for r in items.results() if r.status != 0)
        print(json.dumps({"threads": threads, "inflight": inflight, "images_per_s": round(n / ms * 1e3, 1),
                          "device_ms_per_image": round(ms / n, 4), "host_wall_ms": round(wall, 1), "failed": bad}), flush=True)
