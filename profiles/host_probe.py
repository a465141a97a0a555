"""Where the host thread's time goes per image in the resident-pixel loop (submit vs wait)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import image_compression_b200 as pkg
from image_compression_b200 import synth
w, h, target, q = 4000, 3000, 1048576, float(sys.argv[2]) if len(sys.argv) > 2 else 0.25
slots = int(sys.argv[1]) if len(sys.argv) > 1 else 6
eng = pkg.Engine(devices=[0], slots_per_gpu=slots, max_w=w, max_h=h)
pool = [synth.natural_image(1000 + i, w, h) for i in range(4)]
res = [eng.upload_image(p, 0) for p in pool]
outs = [eng.alloc_pinned((target + 16,)) for _ in range(slots)]
def run(n):
    ts = tw = 0.0
    inflight = []
    t00 = time.perf_counter()
    for i in range(n):
        if len(inflight) == slots:
            t0 = time.perf_counter(); eng.wait(inflight.pop(0)); tw += time.perf_counter() - t0
        t0 = time.perf_counter(); inflight.append(eng.submit_jpg_fit_image(res[i % 4], q, target, outs[i % slots])); ts += time.perf_counter() - t0
    while inflight:
        t0 = time.perf_counter(); eng.wait(inflight.pop(0)); tw += time.perf_counter() - t0
    tot = time.perf_counter() - t00
    return tot / n * 1e6, ts / n * 1e6, tw / n * 1e6
run(64)
for _ in range(3):
    print("slots %d: per image total %.1f us, submit %.1f us, wait %.1f us" % ((slots,) + run(512)))
